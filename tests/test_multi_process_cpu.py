"""CPU, world_size 2 over gloo: the host-side logic of chain sharding (chain ranges, global-chain-keyed draw
streams, gather in chain order).  The per-rank back end here is the CPU oracle; on GPUs the same code runs with the
CUDA library (tests/test_multi_gpu.py)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import nimble_b200 as nb
    from nimble_b200 import parallel
    from oracle.oracle_capi import oracle_backend
    from tests import small_models as sm
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    m, conf = sm.SMALL["pump"]()

    def factory(first_chain, dev):
        return nb.compileNimble(nb.buildMCMC(conf), backend=oracle_backend(), first_chain=first_chain)
    inits = [{"alpha": 1.0 + 0.1 * c} for c in range(5)]
    s = parallel.run_chain_sharded(factory, 120, 20, 1, nchains=5, inits=inits, seed=9)
    if rank == 0:
        np.save(out_path, s)
    dist.barrier()
    dist.destroy_process_group()


def test_hybrid_layout():
    from nimble_b200.parallel import hybrid_layout
    assert [hybrid_layout(r, 8, 4) for r in range(8)] == [(0, 0, 2), (0, 1, 2), (0, 2, 2), (0, 3, 2), (1, 0, 2), (1, 1, 2),
                                                          (1, 2, 2), (1, 3, 2)]
    assert hybrid_layout(3, 4, 1) == (3, 0, 4) and hybrid_layout(3, 4, 4) == (0, 3, 1)
    with pytest.raises(ValueError):
        hybrid_layout(0, 6, 4)


def test_chain_ranges_cover_everything():
    from nimble_b200.parallel import chain_range, node_range
    for n, w in ((5, 2), (1024, 8), (3, 4), (7, 7)):
        spans = [chain_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
    for n, w in ((1_000_000, 8), (1000, 3), (100, 8), (10_000_000, 2)):
        spans = [node_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
        assert all(lo % 128 == 0 or lo == n for lo, _ in spans)


def test_two_ranks_reproduce_the_single_process_run(tmp_path, oracle):
    import nimble_b200 as nb
    from tests import small_models as sm
    out = str(tmp_path / "s.npy")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    sharded = np.load(out)
    m, conf = sm.SMALL["pump"]()
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=oracle)
    inits = np.stack([m.init_vector({"alpha": 1.0 + 0.1 * c}) for c in range(5)])
    single = cm.run_chains(inits, 120, 20, 1, True, 9)
    assert sharded.shape == single.shape == (5, 100, 12)
    assert np.array_equal(sharded, single)

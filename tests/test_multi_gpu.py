"""GPU, two ranks on two GPUs: data-node sharding with the all-reduce inside the library (its own kernel over NVLink peer memory,
or NCCL), and chain sharding.
Skipped unless the box has at least two GPUs -- except test_two_ranks_on_one_device_share_the_exchange, which puts both ranks
of a group on device 0 (their all-reduce kernels are one block each and co-resident; a rank that never arrives times out inside
the kernel) so that the data-sharded path is exercised on a one-GPU box as well."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count()


def _data_shard_worker(rank, world, port, name, out_dir, transport="nccl"):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import nimble_b200 as nb
    from nimble_b200 import _capi
    from tests import small_models as sm
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    backend = _capi.default_backend()
    m, conf = sm.SMALL[name]()
    C = 4
    inits = sm.random_inits(m, C, 21, 0.1)
    if name == "occupancy":
        inits[:, m.var_slot["z"]:m.var_slot["z"] + 300] = 1.0
    from nimble_b200 import parallel
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=rank)
    dm = cm._ensure(C)
    parallel.attach_data_sharding(dm, world, rank, transport=transport, min_nodes=256)
    s = cm.run_chains(inits, 260, 10, 1, True, 17)
    np.save(os.path.join(out_dir, "rank%d.npy" % rank), s)
    n_ar, replayed = dm.last_allreduces()
    assert n_ar > 0
    if transport == "peer":
        assert replayed, "the peer all-reduce is an ordinary kernel: the iteration must have been replayed from its CUDA graph"
    if rank == 0:
        ref = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=0)
        np.save(os.path.join(out_dir, "single.npy"), ref.run_chains(inits, 260, 10, 1, True, 17))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("transport", ["peer", "nccl"])
@pytest.mark.parametrize("name", ["logistic", "lmm", "glmm", "occupancy"])
def test_data_sharded_run_matches_single_gpu(tmp_path, name, transport):
    if _ngpu() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    port = 29600 + os.getpid() % 2000
    mp.spawn(_data_shard_worker, args=(2, port, name, str(tmp_path), transport), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    single = np.load(tmp_path / "single.npy")
    assert np.array_equal(r0, r1)                    # every rank took identical decisions
    err = np.abs(r0 - single) / np.maximum(1.0, np.abs(single))
    assert err.max() < 1e-9, err.max()               # same chains as the unsharded run (summation order differs)


def _hybrid_worker(rank, world, port, name, data_shards, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import nimble_b200 as nb
    from nimble_b200 import _capi, parallel
    from tests import small_models as sm
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    backend = _capi.default_backend()
    m, conf = sm.SMALL[name]()
    C = 6
    inits = sm.random_inits(m, C, 23, 0.1)

    def factory(first_chain, dev):
        return nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=dev, first_chain=first_chain)

    # run_hybrid flattens per-chain inits through model.init_vector: let ready-made state vectors pass through
    orig = m.init_vector
    m.init_vector = lambda these=None: orig() if these is None else np.asarray(these, dtype=np.float64)
    s = parallel.run_hybrid(factory, 200, 10, 1, nchains=C, inits=[inits[c] for c in range(C)], seed=19,
                            data_shards=data_shards, min_nodes=256)
    if rank == 0:
        np.save(os.path.join(out_dir, "hybrid.npy"), s)
        ref = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=0)
        np.save(os.path.join(out_dir, "single.npy"), ref.run_chains(inits, 200, 10, 1, True, 19))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,data_shards", [(2, 1), (2, 2), (4, 2)])
def test_hybrid_chain_groups_times_data_shards(tmp_path, world, data_shards):
    """SURVEY.md 8e third row: chain groups x data shards, one NCCL communicator per group; (2,1) is plain chain sharding,
    (2,2) plain data sharding, (4,2) two groups of two shards."""
    if _ngpu() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    port = 29600 + (os.getpid() + 7 * world + data_shards) % 2000
    mp.spawn(_hybrid_worker, args=(world, port, "lmm", data_shards, str(tmp_path)), nprocs=world, join=True)
    h, single = np.load(tmp_path / "hybrid.npy"), np.load(tmp_path / "single.npy")
    assert h.shape == single.shape
    err = np.abs(h - single) / np.maximum(1.0, np.abs(single))
    assert err.max() < 1e-9, err.max()


def test_two_devices_in_one_process_share_the_exchange():
    """Two models on two devices of ONE process (per-device kernel attributes, the same-process branch of the peer exchange):
    each is driven from its own host thread, because a run blocks its caller until the device is done and the two ranks wait
    for one another on the device."""
    if _ngpu() < 2:
        pytest.skip("needs two GPUs")
    import threading
    import nimble_b200 as nb
    from nimble_b200 import _capi
    from tests import small_models as sm
    backend = _capi.default_backend()
    m, conf = sm.SMALL["logistic"]()
    C = 4
    inits = sm.random_inits(m, C, 21, 0.1)
    cms = [nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=d) for d in (0, 1)]
    dms = [cm._ensure(C) for cm in cms]
    handles = [dm.peer_export(2) for dm in dms]
    for r, dm in enumerate(dms):
        dm.enable_data_sharding_peer(2, r, handles, min_nodes=256)
    out, errs = [None, None], []

    def work(r):
        try:
            out[r] = cms[r].run_chains(inits, 120, 10, 1, True, 17)
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)
    ths = [threading.Thread(target=work, args=(r,)) for r in (0, 1)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    assert np.array_equal(out[0], out[1])
    single = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=1).run_chains(inits, 120, 10, 1, True, 17)  # fused kernel on device 1
    err = np.abs(out[0] - single) / np.maximum(1.0, np.abs(single))
    assert err.max() < 1e-9, err.max()


@pytest.mark.parametrize("name", ["logistic", "lmm", "glmm", "occupancy"])
def test_two_ranks_on_one_device_share_the_exchange(name):
    """The data-sharded path on a one-GPU box (tests/one_device_ranks.py): two models on device 0 of ONE process are the two
    ranks of a group, each driven from its own host thread.  Run in a child process with eager module loading: loading a
    kernel lazily synchronises the context, which cannot complete while the other rank's all-reduce kernel spins."""
    if _ngpu() < 1:
        pytest.skip("needs a GPU")
    import subprocess
    import sys
    env = dict(os.environ, CUDA_MODULE_LOADING="EAGER")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "tests.one_device_ranks", name], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])

"""bench.py helper for --gpus N > 1: the data-sharded configuration of BASELINE.json (configs[2], the linear mixed model:
8 chains, every rank holds all chains and a contiguous slice of the N data nodes, each per-proposal block sum all-reduced
inside the device library).  Reports, next to the chain-sharded headline:
  parity      a scaled-down run on all ranks against rank 0's unsharded run of the same chains (max relative difference),
              and that every rank produced the same bits;
  per size    ms per MCMC iteration data-sharded over N GPUs (device-timed, max over ranks) and on ONE GPU, all-reduces per
              iteration, the latency of one all-reduce alone, its share of the iteration, and whether the iteration ran as a
              replayed CUDA graph -- for the library's own peer-memory all-reduce and for host-issued NCCL."""
import time

import numpy as np


def _ar_max(torch, dist, x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _sharded_run(nb, backend, conf, inits, local, rank, world, transport, niter, warm, min_nodes, torch, dist, barrier):
    from nimble_b200 import parallel
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local)
    dm = cm._ensure(inits.shape[0])
    parallel.attach_data_sharding(dm, world, rank, transport=transport, min_nodes=min_nodes)
    dm.set_values(inits)
    dm.run(warm, 0, 1, True, 31)
    barrier()
    dm.run(niter, 0, 1, False, 31)
    barrier()
    ms, launches = dm.last_timing()
    n_ar, replayed = dm.last_allreduces()
    ar_ms = dm.time_allreduce(inits.shape[0], 200)
    out = {"ms_per_iteration": _ar_max(torch, dist, ms) / niter, "allreduces_per_iteration": n_ar / niter,
           "allreduce_us": _ar_max(torch, dist, ar_ms) * 1e3, "graph_replayed": bool(replayed),
           "launches_per_iteration": launches / niter}
    out["allreduce_share_of_iteration"] = out["allreduces_per_iteration"] * out["allreduce_us"] * 1e-3 / out["ms_per_iteration"]
    samples = dm.samples()
    dm.close()
    return out, samples


def data_sharded_block(a, nb, backend, local, rank, world, dist, torch, barrier):
    from nimble_b200 import configs
    C = 8
    res = {"what": "configs[2] linear mixed model, %d chains, data nodes sharded over %d GPUs (all chains on every GPU); all-reduce "
                   "of `chains` doubles per block sum" % (C, world)}
    # ---- parity inside this invocation: small model, sharded on all ranks vs unsharded on rank 0
    m, conf = configs.lmm(N=64 * 1024, G=64, seed=1003)
    inits = np.tile(m.init_vector(), (C, 1))
    inits[:, m.var_slot["mu"]] = np.linspace(1.0, 3.0, C)
    _, s_sh = _sharded_run(nb, backend, conf, inits, local, rank, world, "peer", 150, 10, 1024, torch, dist, barrier)
    digest = torch.tensor([float(np.sum(s_sh)), float(np.sum(s_sh * s_sh))], dtype=torch.float64, device="cuda")
    lo, hi = digest.clone(), digest.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    par = {"ranks_bit_identical": bool(torch.equal(lo, hi)), "model": "lmm N=65536 G=64, %d chains, 160 iterations" % C}
    if rank == 0:
        ref = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local)
        d1 = ref._ensure(C)
        d1.set_values(inits)
        d1.run(10, 0, 1, True, 31)
        d1.run(150, 0, 1, False, 31)
        s1 = d1.samples()
        d1.close()
        par["max_rel_diff_vs_unsharded"] = float(np.max(np.abs(s_sh - s1) / np.maximum(1.0, np.abs(s1))))
        par["pass"] = par["ranks_bit_identical"] and par["max_rel_diff_vs_unsharded"] < 1e-9
    res["parity"] = par
    barrier()
    # ---- timing at full size and at a size where sharding should pay
    sizes = []
    for N in (10_000_000, 80_000_000):
        t0 = time.time()
        m, conf = configs.lmm(N=N, G=10_000, seed=1003)
        inits = np.tile(m.init_vector(), (C, 1))
        entry = {"N": N, "host_build_seconds": time.time() - t0}
        for transport in ("peer", "nccl"):
            r, _ = _sharded_run(nb, backend, conf, inits, local, rank, world, transport, 200, 20, 1 << 16, torch, dist, barrier)
            entry[transport] = r
        if rank == 0:   # the same model on one GPU, the other ranks wait
            cm1 = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local)
            d1 = cm1._ensure(C)
            d1.set_values(inits)
            d1.run(20, 0, 1, True, 31)
            d1.run(200, 0, 1, False, 31)
            ms1, l1 = d1.last_timing()
            d1.close()
            entry["one_gpu_ms_per_iteration"] = ms1 / 200
            entry["speedup_peer_vs_one_gpu"] = entry["one_gpu_ms_per_iteration"] / entry["peer"]["ms_per_iteration"]
            entry["speedup_nccl_vs_one_gpu"] = entry["one_gpu_ms_per_iteration"] / entry["nccl"]["ms_per_iteration"]
        barrier()
        sizes.append(entry)
        del m, conf
    res["sizes"] = sizes
    return res

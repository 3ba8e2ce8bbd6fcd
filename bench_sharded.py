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
    # ---- timing.  Cases: (1) the LMM as north_star names it, every block sum a streaming pass over this rank's slice plus an
    # all-reduce (group-statistics shortcut switched off); (2) the LMM as the library runs it by default (block sums from data-only
    # group statistics, all-reduced once: nothing left to exchange per iteration); (3) the same streaming layout at N = 8e7, where
    # a slice is still 10x the L2; (4) configs[1] with few chains (64 in total, all on every GPU): the fused logistic kernel over
    # each rank's tiles of X; (5) configs[4] at 8192 chains with group-aligned shards.
    import os
    cases = []

    def case(label, make, C_, niter, warm, env=None, transports=("peer", "nccl")):
        t0 = time.time()
        for k, v in (env or {}).items():
            os.environ[k] = v
        try:
            m_, conf_, inits_ = make(C_)
            entry = {"case": label, "chains": C_, "host_build_seconds": time.time() - t0}
            for transport in transports:
                r, _ = _sharded_run(nb, backend, conf_, inits_, local, rank, world, transport, niter, warm, 1 << 16, torch, dist, barrier)
                entry[transport] = r
            if rank == 0:   # the same model on one GPU, the other ranks wait
                cm1 = nb.compileNimble(nb.buildMCMC(conf_), backend=backend, device=local)
                d1 = cm1._ensure(C_)
                d1.set_values(inits_)
                d1.run(warm, 0, 1, True, 31)
                d1.run(niter, 0, 1, False, 31)
                ms1, _ = d1.last_timing()
                d1.close()
                entry["one_gpu_ms_per_iteration"] = ms1 / niter
                for transport in transports:
                    entry["speedup_%s_vs_one_gpu" % transport] = entry["one_gpu_ms_per_iteration"] / entry[transport]["ms_per_iteration"]
            barrier()
        except Exception as ex:  # noqa: BLE001 -- one case must not take the others down
            entry = {"case": label, "error": "%s: %s" % (type(ex).__name__, str(ex)[:200])}
        finally:
            for k in (env or {}):
                os.environ.pop(k, None)
        cases.append(entry)

    def lmm_of(N):
        def make(C_):
            m_, conf_ = configs.lmm(N=N, G=10_000, seed=1003)
            return m_, conf_, np.tile(m_.init_vector(), (C_, 1))
        return make

    def logistic_few(C_):
        m_, conf_ = configs.logistic(N=a.n, p=a.p, seed=1001)
        return m_, conf_, np.stack([m_.init_vector(i) for i in configs.logistic_inits(a.p, C_)])

    def glmm_full(C_):
        m_, conf_ = configs.glmm()
        return m_, conf_, np.tile(m_.init_vector(), (C_, 1))

    stream = {"NIMB200_NO_BYPRODUCT": "1"}
    case("lmm N=1e7, streaming block sums + all-reduce (NIMB200_NO_BYPRODUCT)", lmm_of(10_000_000), C, 200, 20, stream)
    case("lmm N=1e7, default (group statistics)", lmm_of(10_000_000), C, 200, 20, None, ("peer",))
    case("lmm N=8e7, streaming block sums + all-reduce (NIMB200_NO_BYPRODUCT)", lmm_of(80_000_000), C, 100, 10, stream, ("peer",))
    case("logistic N=%d p=%d, 64 chains in total on every GPU" % (a.n, a.p), logistic_few, 64, 100, 10)
    case("glmm N=1e6 G=1e4, 8192 chains, group-aligned shards", glmm_full, 8192, 20, 3, None, ("peer",))
    res["cases"] = cases
    return res

"""bench.py helper: live rooflines of the family kernels behind BASELINE.json configs[2..4], each on its own configuration at
full size (rank 0, single-GPU runs): one block sum of the data declaration, CUDA events, min of 3 x reps launches."""
import numpy as np


def _time(dm, decl, reps):
    return min(dm.time_calculate(decl, reps) for _ in range(3))


def secondary_rooflines(nb, backend, local, hbm_peak, dfma_tflops):
    from nimble_b200 import configs
    out = []
    try:   # configs[2]: LMM, N = 1e7, 8 chains -- the one HBM-bound logProb kernel of the path
        m3, conf3 = configs.lmm()
        cm = nb.compileNimble(nb.buildMCMC(conf3), backend=backend, device=local)
        dl = cm._ensure(8)
        dl.set_values(np.tile(m3.init_vector(), (8, 1)))
        yd = len(cm.builder.decls) - 1
        nn = cm.builder.decls[yd].n
        ms = _time(dl, yd, 50)
        by = 12.0 * nn
        out.append({"kernel": "k_norm_stream_tma + k_norm_reduce_finish", "bound": "hbm",
                    "workload": "configs[2]: y[i] ~ dnorm(mu + u[grp[i]], sd) block sum, N = %d, 8 chains" % nn,
                    "achieved": by / (ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": by / (ms * 1e-3) / 1e9 / hbm_peak,
                    "us_per_block_sum": ms * 1e3, "algorithmic_bytes": by,
                    "note": "both launches of one block sum are inside the timed region"})
        dl.close()
    except Exception as ex:  # the headline line must not depend on it
        out.append({"kernel": "k_norm_stream_tma", "error": str(ex)[:200]})
    try:   # configs[4]: Poisson GLMM, N = 1e6, 8192 chains -- FP64-instruction-bound
        m5, conf5 = configs.glmm()
        cm = nb.compileNimble(nb.buildMCMC(conf5), backend=backend, device=local)
        C = 8192
        dg = cm._ensure(C)
        init5 = np.tile(m5.init_vector(), (C, 1))
        rng = np.random.default_rng(5)
        init5[:, m5.var_slot["b0"]] = 0.5 + 0.1 * rng.standard_normal(C)
        init5[:, m5.var_slot["b1"]] = 0.3 + 0.05 * rng.standard_normal(C)
        init5[:, m5.var_slot["lam"]:m5.var_slot["lam"] + 10_000] = rng.gamma(2.0, 0.5, (C, 10_000))
        dg.set_values(init5)
        yd = len(cm.builder.decls) - 1
        nn = cm.builder.decls[yd].n
        ms_groups = _time(dg, yd, 5)     # default path: the block sum from per-group sums S_g(b1) (no node is read once S is in place)
        import os
        os.environ["NIMB200_NO_BYPRODUCT"] = "1"   # the node kernel itself
        try:
            ms = _time(dg, yd, 5)
        finally:
            del os.environ["NIMB200_NO_BYPRODUCT"]
        fp64_per_eval = 16.0     # FP64-pipe instructions per evaluation as implemented (ncu-counted, profiles/r2l_pois_sum_raw.csv)
        ach = 2.0 * fp64_per_eval * nn * C / (ms * 1e-3) / 1e12
        out.append({"kernel": "k_pois_sum", "bound": "fp64",
                    "workload": "configs[4]: y[i] ~ dpois(lam[grp[i]] exp(b0 + b1 x[i])) block sum, N = %d, %d chains" % (nn, C),
                    "achieved": ach, "peak": dfma_tflops, "unit": "TFLOP/s", "frac": ach / dfma_tflops if dfma_tflops else None,
                    "ms_per_block_sum": ms, "evals_per_s": nn * C / (ms * 1e-3), "fp64_instructions_per_eval": fp64_per_eval,
                    "ms_per_block_sum_from_group_sums": ms_groups,
                    "note": "an MCMC iteration no longer runs this kernel: block sums and the conjugate sweep come from per-group sums "
                            "S_g = sum exp(b1 x), themselves evaluated from 25 data-only moment coefficients per group (24 fused "
                            "multiply-adds per group and chain) while |b1| max|x| <= 2, by a pass over the nodes otherwise",
                    "what": "FP64-pipe instructions the kernel issues per evaluation (2 flop each) against the DFMA peak measured live"})
        dg.close()
    except Exception as ex:
        out.append({"kernel": "k_pois_sum", "error": str(ex)[:200]})
    try:   # configs[3]: occupancy, S = 1e5 x J = 10, 64 chains
        m4, conf4 = configs.occupancy()
        cm = nb.compileNimble(nb.buildMCMC(conf4), backend=backend, device=local)
        C = 64
        do = cm._ensure(C)
        init4 = np.tile(m4.init_vector(), (C, 1))
        rng = np.random.default_rng(4)
        for k, v in (("b0", 0.3), ("b1", 1.0), ("a0", -0.5), ("a1", 0.7)):   # chains spread around the generator's values
            init4[:, m4.var_slot[k]] = v + 0.1 * rng.standard_normal(C)
        do.set_values(init4)
        yd = [cm.lowering.decl_ids[d.index] for d in m4.decls if d.var == "y"][0]
        nn = cm.builder.decls[yd].n
        ms = _time(do, yd, 20)
        fp64_per_eval = 23.0
        ach = 2.0 * fp64_per_eval * nn * C / (ms * 1e-3) / 1e12
        out.append({"kernel": "k_bernlin_sum", "bound": "fp64",
                    "workload": "configs[3]: y[i,j] ~ dbern(z[i] ilogit(a0 + a1 w[i,j])) block sum, %d nodes, %d chains" % (nn, C),
                    "achieved": ach, "peak": dfma_tflops, "unit": "TFLOP/s", "frac": ach / dfma_tflops if dfma_tflops else None,
                    "us_per_block_sum": ms * 1e3, "evals_per_s": nn * C / (ms * 1e-3), "fp64_instructions_per_eval": fp64_per_eval,
                    "what": "FP64-pipe instructions of the reciprocal-free dbern-logit form per evaluation (2 flop each) against the "
                            "DFMA peak measured live; the kernel is issue-bound (57 other instructions per evaluation, each costing "
                            "about a cycle on top of two per FP64 instruction), see DESIGN.md"})
        do.close()
    except Exception as ex:
        out.append({"kernel": "k_bernlin_sum", "error": str(ex)[:200]})
    try:   # the issue model behind the instruction budgets: DFMA rate with 0 / 4 / 8 / 16 integer instructions per 8 DFMAs
        mix = backend.fp64_issue_mix(local)
        out.append({"kernel": "k_dfma_int_mix (microbenchmark)", "unit": "TFLOP/s", "dfma_tflops_with_0_4_8_16_int_per_8_dfma": [float(v) for v in mix]})
    except Exception as ex:
        out.append({"kernel": "k_dfma_int_mix", "error": str(ex)[:200]})
    return out

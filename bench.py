#!/usr/bin/env python3
"""bench.py -- headline measurement for the B200 multi-chain MCMC back end.

Workload (BASELINE.json configs[1]): hierarchical logistic regression, N = 1e6 observations, p = 50,
RW_block on beta[1:50] + RW(log) on tau_b + conjugate mu_b, 1024 chains PER GPU (chains are sharded across
ranks with no collective: "scaling": "weak").  One step = one MCMC iteration of every chain on the rank.

  python bench.py --gpus N --steps K --warmup W          (N > 1: launched by torch.distributed.run)
  python bench.py --impl reference ...                   the CPU oracle (restatement of compiled NIMBLE; R and
                                                         Rmath are not installable here) on all host cores.
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "logprob_evals_per_sec"
UNIT = "logProb evals/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--chains", type=int, default=1024, help="chains per GPU")
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--p", type=int, default=50)
    ap.add_argument("--cpu-chains", type=int, default=0, help="chains of the CPU baseline (default: host cores, <= 64)")
    ap.add_argument("--cpu-iters", type=int, default=12)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the live measurement of the HBM-bound kernel of configs[2] (rank 0, single-GPU runs only)")
    ap.add_argument("--ess-niter", type=int, default=5000,
                    help="iterations of the min-ESS/s protocol run (SURVEY 8d: niter 5000, nburnin = niter / 2); 0 skips it")
    ap.add_argument("--ess-chains", type=int, default=0, help="chains the ESS estimate uses (0 = all chains of the rank)")
    ap.add_argument("--ref-ess-budget-s", type=float, default=240.0,
                    help="reference arm: wall-clock budget of its min-ESS/s protocol run; it stops early (and says so) beyond it")
    return ap.parse_args()


def workload_config(a, extra=None):
    c = {"workload": "hierarchical logistic regression N=%d p=%d, RW_block(beta[1:%d]) + RW(log tau_b) + conjugate mu_b"
         % (a.n, a.p, a.p), "chains_per_gpu": a.chains, "step": "one MCMC iteration of every chain",
         "l2": "inputs larger than L2 (X is %.0f MB)" % (a.n * a.p * 8 / 1e6), "parallelism": "chain-sharded, no collective"}
    if extra:
        c.update(extra)
    return c


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_model(a):
    from nimble_b200 import configs
    t0 = time.time()
    m, conf = configs.logistic(N=a.n, p=a.p, seed=1001)
    return m, conf, time.time() - t0


def _oracle_runner(a, conf, model, chains, seed):
    """The oracle (CPU restatement of compiled NIMBLE's sampling path) with one chain per host thread, the
    one-process-per-core pattern of the NIMBLE manual (chapter_MCMC.Rmd:1223-1229).  Returns (dm, run, evals_per_iteration)."""
    import nimble_b200 as nb
    from nimble_b200 import configs
    from oracle.oracle_capi import oracle_backend
    ob = oracle_backend()
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=ob)
    dm = cm._ensure(chains)
    inits = np.stack([model.init_vector(i) for i in configs.logistic_inits(a.p, chains)])
    dm.set_values(inits)

    def run(n_it, reset):
        ob.check(ob.prepare(dm.h, n_it, 0, 1, 1 if reset else 0))
        ths = [threading.Thread(target=ob.run_range, args=(dm.h, n_it, 0, 1, 1 if reset else 0, seed, c, c + 1))
               for c in range(chains)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        ob.check(ob.finish(dm.h, n_it, 0, 1))
        return dt
    return dm, run, configs.evals_per_iteration(cm), cm


def cpu_baseline(a, conf, model, niter, warm=1):
    cores = os.cpu_count() or 1
    chains = a.cpu_chains or min(cores, 64)
    dm, run, evals_iter, _ = _oracle_runner(a, conf, model, chains, 12345)
    if warm:
        run(warm, True)
    dt = run(niter, False if warm else True)
    val = evals_iter * chains * niter / dt
    return {"value": val, "unit": UNIT, "cores": min(chains, cores), "kind": "port",
            "sample": "%d chains x %d iterations of the same N=%d p=%d workload, one chain per host thread (%d cores)"
                      % (chains, niter, a.n, a.p, cores),
            "ms_per_iteration_per_chain": dt / niter * 1e3, "seconds": dt}, dt, chains


ESS_SEED = 777


def cpu_min_ess(a, conf, model):
    """min-ESS/s of the reference arm from ITS OWN samples (MCMC efficiency as P/R/MCMC_autoBlock.R:297-306 defines it: min over
    the monitored columns of effectiveSize / run time; sums over chains, run time includes burn-in).  Same protocol, seed and
    initial values as the GPU arm's min_ess block, so these chains are the first `chains` chains of that run.  The run proceeds
    in segments and stops early once the wall-clock budget is spent (the line then says how far it got)."""
    from nimble_b200.ess import min_ess
    cores = os.cpu_count() or 1
    chains = a.cpu_chains or min(cores, 64)
    dm, run, evals_iter, cm = _oracle_runner(a, conf, model, chains, ESS_SEED)
    seg, done, wall = 250, 0, 0.0
    while done < a.ess_niter:
        n_it = min(seg, a.ess_niter - done)
        if done and wall + wall / done * n_it > a.ref_ess_budget_s:
            break
        wall += run(n_it, done == 0)
        done += n_it
    s = dm.samples()[:, done // 2:done, :]
    mess, tot = min_ess(s)
    return {"min_ess_per_sec": mess / wall if wall > 0 else None, "min_ess_total": mess, "chains": chains, "niter": done,
            "nburnin": done // 2, "protocol_niter": a.ess_niter, "complete": done == a.ess_niter, "seconds": wall,
            "worst_column": cm.colnames[int(np.argmin(np.where(tot > 0, tot, np.inf)))] if np.any(tot > 0) else None,
            "zero_ess_columns": int((tot <= 0).sum()), "seed": ESS_SEED, "cores": min(chains, cores),
            "definition": "min over monitored columns of (sum over chains of coda-style effectiveSize of the post-burn-in "
                          "samples) / wall seconds of the whole run incl. burn-in (MCMC_autoBlock.R:297-306)"}


def _emit(line):
    """The JSON line goes to the process's ORIGINAL stdout; everything else that lands on fd 1 meanwhile (NCCL's version
    banner, library chatter) has been diverted to stderr, so stdout carries exactly one line."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def reference_arm(a):
    model, conf, _ = build_model(a)
    cb, dt, chains = cpu_baseline(a, conf, model, max(1, a.steps), warm=max(1, min(a.warmup, 3)))
    ess = None
    if a.ess_niter > 0 and a.gpus == 1:
        ess = cpu_min_ess(a, conf, model)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": dt / max(1, a.steps) * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(a, {"chains": chains, "note": "reference's compiled NIMBLE cannot be built "
                                          "here (no R / Rmath); this is its CPU restatement (oracle/), one chain per core"}),
            "cpu_baseline": cb,
            "min_ess_per_sec": ess["min_ess_per_sec"] if ess else None,
            "min_ess": ess if ess else ("not run for --gpus > 1 (the CPU arm does not depend on the GPU count): see the --gpus 1 line"
                                        if a.ess_niter > 0 else "skipped (--ess-niter 0)"),
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(line)
    return 0


def gpu_min_ess(a, nb, conf, backend, local, first_chain, inits, C, world, dist, torch):
    """The SURVEY 8d protocol on this rank's chains: niter, nburnin = niter / 2, every chain; min-ESS/s by the reference's
    definition of MCMC efficiency (MCMC_autoBlock.R:297-306).  ESS sums over all chains of all ranks, time is the slowest
    rank's wall clock around the run (burn-in included, sample read-back excluded, as the reference times mcmc$run)."""
    from nimble_b200.ess import min_ess
    cm3 = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=first_chain)
    d3 = cm3._ensure(C)
    d3.set_values(inits)
    nb_it = a.ess_niter // 2
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    d3.run(a.ess_niter, nb_it, 1, True, ESS_SEED)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ms_dev, _ = d3.last_timing()
    sf = d3.samples()
    d3.close()
    nes = min(a.ess_chains or C, C)
    t1 = time.perf_counter()
    _, tot = min_ess(sf[:nes])
    t_ess = time.perf_counter() - t1
    tot = tot * (C / nes)
    tt = torch.tensor(np.concatenate([tot, [0.0]]), dtype=torch.float64, device="cuda")
    tw = torch.tensor([wall, ms_dev * 1e-3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
    tot = tt.cpu().numpy()[:-1]
    wall, dev_s = (float(x) for x in tw.cpu().numpy())
    nz = tot[tot > 0]
    mess = float(nz.min()) if nz.size else 0.0
    pooled = sf.reshape(-1, sf.shape[-1])
    return {"min_ess_per_sec": mess / wall, "min_ess_total": mess, "chains": C * world, "niter": a.ess_niter, "nburnin": nb_it,
            "seconds": wall, "device_seconds": dev_s, "ess_estimated_on_chains_per_rank": nes, "ess_host_seconds": t_ess,
            "worst_column": cm3.colnames[int(np.argmin(np.where(tot > 0, tot, np.inf)))] if nz.size else None,
            "zero_ess_columns": int((tot <= 0).sum()), "seed": ESS_SEED,
            "posterior_mean_first_cols": dict(zip(cm3.colnames[:4], [float(x) for x in pooled.mean(0)[:4]])),
            "posterior_sd_first_cols": dict(zip(cm3.colnames[:4], [float(x) for x in pooled.std(0)[:4]])),
            "definition": "min over monitored columns of (sum over chains of coda-style effectiveSize of the post-burn-in "
                          "samples) / wall seconds of the whole run incl. burn-in (MCMC_autoBlock.R:297-306)"}


def strong_scaling_block(a, nb, conf, model, backend, local, rank, world, dist, torch, barrier, total_chains=1024):
    """BASELINE.json configs[1] as north_star states it for the 8-GPU box: 1024 chains IN TOTAL, chain-sharded (1024 / N per
    GPU, no collective).  Same K timed steps after W warm-up steps, device-timed, max over ranks."""
    from nimble_b200 import configs
    from nimble_b200.parallel import chain_range
    lo, hi = chain_range(total_chains, rank, world)
    Cs = hi - lo
    inits_all = configs.logistic_inits(a.p, total_chains)
    inits = np.stack([model.init_vector(i) for i in inits_all[lo:hi]])
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=lo)
    evals_iter = configs.evals_per_iteration(cm)
    dm = cm._ensure(Cs)
    dm.set_values(inits)
    dm.run(a.warmup, 0, 1, True, 2024)
    barrier()
    dm.run(a.steps, 0, 1, False, 2024)
    barrier()
    ms, launches = dm.last_timing()
    ydecl = len(cm.builder.decls) - 1
    k_ms = dm.time_calculate(ydecl, 10)
    dm.close()
    t = torch.tensor([ms, k_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, k_ms = (float(x) for x in t.cpu().numpy())
    out = {"what": "strong scaling: %d chains in total, %d per GPU, chain-sharded, no collective" % (total_chains, Cs),
           "chains_total": total_chains, "chains_per_gpu": Cs, "ms_per_step": ms / a.steps,
           "value": evals_iter * total_chains * a.steps / (ms * 1e-3), "unit": UNIT,
           "fused_kernel_ms": k_ms, "non_kernel_ms_per_step": ms / a.steps - k_ms, "launches": int(launches)}
    if a.ess_niter > 0:   # the north-star's own figure: min-ESS/s of the 1024-chain job on N GPUs (same protocol and seed)
        e = gpu_min_ess(a, nb, conf, backend, local, lo, inits, Cs, world, dist, torch)
        out["min_ess_per_sec"] = e["min_ess_per_sec"]
        out["min_ess"] = {k: e[k] for k in ("min_ess_total", "chains", "niter", "nburnin", "seconds", "worst_column")}
    return out


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if a.impl == "reference":
        return reference_arm(a) if rank == 0 else 0

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the B200 back end has no CPU fallback"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    import nimble_b200 as nb
    from nimble_b200 import _capi, configs

    backend = _capi.default_backend()
    model, conf, t_build = build_model(a)
    C = a.chains
    first_chain = rank * C
    inits_all = configs.logistic_inits(a.p, C * world)
    inits = np.stack([model.init_vector(i) for i in inits_all[first_chain:first_chain + C]])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput: inputs already in HBM, K timed iterations, CUDA events on the run stream
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=first_chain)
    evals_iter = configs.evals_per_iteration(cm)
    dm = cm._ensure(C)
    dm.set_values(inits)
    dm.run(a.warmup, 0, 1, True, 2024)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    dm.run(a.steps, 0, 1, False, 2024)
    barrier()
    clk = clocks.stop()
    ms, launches = dm.last_timing()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = evals_iter * C * world * a.steps / (ms_max * 1e-3)

    # ---- end to end through the public API with host buffers: compile (lowering) once, then ONE runMCMC call that
    # uploads the table (X, y, ...) and the inits, runs W+K iterations and reads mvSamples back.
    cm2 = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=first_chain)
    barrier()
    t0 = time.perf_counter()
    nb.runMCMC(cm2, niter=a.warmup + a.steps, nburnin=a.warmup, nchains=C, inits=inits_all[first_chain:first_chain + C],
               setSeed=2024)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    te = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    t_e2e = float(te.item())
    b = cm2.builder
    h2d = sum(x.nbytes for x in b.darr) + sum(x.nbytes for x in b.iarr) + inits.nbytes
    d2h = C * a.steps * len(b.monitors) * 8
    nsteps_e2e = a.warmup + a.steps
    e2e_val = evals_iter * C * world * nsteps_e2e / t_e2e
    cm2.dm.close()

    # ---- roofline inputs of the dominant kernel (measured on every rank's own device; rank 0's are reported)
    ydecl = len(cm.builder.decls) - 1
    k_ms = dm.time_calculate(ydecl, 10)

    # ---- min-ESS/s, the first half of BASELINE.json's metric (every rank runs the protocol on its own chains)
    ess = gpu_min_ess(a, nb, conf, backend, local, first_chain, inits, C, world, dist, torch) if a.ess_niter > 0 else None

    # ---- N > 1: the north-star's own multi-GPU configurations next to the weak-scaling headline
    strong = sharded = None
    if world > 1:
        dm.close()
        strong = strong_scaling_block(a, nb, conf, model, backend, local, rank, world, dist, torch, barrier)
        try:
            from bench_sharded import data_sharded_block
            sharded = data_sharded_block(a, nb, backend, local, rank, world, dist, torch, barrier)
        except Exception as ex:  # the headline line must not depend on it
            sharded = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300])}

    if rank == 0:
        dfma, dmma = backend.fp64_peaks(local)
        mixed = backend.fp64_mixed_peak(local)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        alg_bytes = 8.0 * a.n * a.p + 8.0 * a.n + 8.0 * a.p * C
        contraction_flops = 2.0 * a.n * a.p * C
        # FP64 work of the link + Bernoulli density per evaluation as the kernel implements it: 24.8
        # FP64-pipe instructions on the reciprocal-free path every realistic linear predictor takes (ncu-counted); one
        # instruction = one multiplier slot = 2 flop in the units of the FMA peak.  Reported next to the strict figure.
        epi_flops = 2.0 * 24.8 * a.n * C
        ach = contraction_flops / (k_ms * 1e-3) / 1e12
        quoted = (a.n, a.p, C) == (1_000_000, 50, 1024)
        roof = {"bound": "tensor", "kernel": "k_fused_bern_logit", "achieved": ach, "peak": dmma, "unit": "TFLOP/s",
                "frac": ach / dmma if dmma else None,
                "peak_source": "FP64 mma.sync loop measured live (nimb200_measure_fp64_peaks); MEASURED_PEAKS.json has no FP64 "
                               "figure. DFMA loop: %.1f TFLOP/s; DMMA+DFMA issued together: %.1f TFLOP/s (the two share the FP64 "
                               "multipliers)" % (dfma, mixed),
                "what": "strict algorithmic flops only: the 2*N*p*C contraction",
                # dram__bytes_read.sum + dram__bytes_write.sum per launch cannot be measured inside this run: the figure is QUOTED
                # from the committed ncu --set full capture of this kernel at this workload, null for any other workload
                "traffic": 429.7e6 if quoted else None,
                "traffic_kind": "quoted from the ncu capture in profiles/ (not measured by this run)" if quoted else None,
                "traffic_source": "profiles/r2f_fused_raw.csv (dram__bytes_read.sum 424.5 MB + dram__bytes_write.sum 5.2 MB)" if quoted else None,
                "algorithmic_bytes": alg_bytes, "kernel_ms": k_ms, "evals_per_s_kernel": a.n * C / (k_ms * 1e-3),
                "fp64_units": {"achieved": (contraction_flops + epi_flops) / (k_ms * 1e-3) / 1e12, "peak": mixed, "unit": "TFLOP/s",
                               "frac": (contraction_flops + epi_flops) / (k_ms * 1e-3) / 1e12 / mixed if mixed else None,
                               "what": "contraction + 24.8 FP64 instructions per evaluation of link and density (as implemented, ncu-counted), "
                                       "against the measured DMMA+DFMA mixed peak"},
                "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                "kernel_share_of_step": k_ms / (ms_max / a.steps)}
        cb = None
        if world == 1 and not a.no_cpu_baseline:
            cb, _, _ = cpu_baseline(a, conf, model, a.cpu_iters)
        secondary = None
        if world == 1 and not a.no_secondary:
            dm.close()
            from bench_secondary import secondary_rooflines
            secondary = secondary_rooflines(nb, backend, local, hbm_peak, dfma)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": workload_config(a, {"chains_total": C * world}),
                "clocks": clk, "gpu_launches": int(launches),
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d / nsteps_e2e, "d2h_bytes_per_step": d2h / nsteps_e2e,
                        "seconds": t_e2e, "steps": nsteps_e2e,
                        "what": "one runMCMC() call on host buffers: table upload + inits + iterations + mvSamples read-back"},
                "roofline": roof, "cpu_baseline": cb,
                "min_ess_per_sec": ess["min_ess_per_sec"] if ess else None, "min_ess": ess,
                "evals_per_iteration_per_chain": evals_iter, "host_lowering_seconds": t_build,
                "secondary_rooflines": secondary, "strong_scaling": strong, "data_sharded": sharded}
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

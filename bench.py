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
    ap.add_argument("--full", action="store_true",
                    help="also run the full protocol of SURVEY 8d (niter 5000, nburnin 2500) and report min-ESS/s from it")
    ap.add_argument("--full-niter", type=int, default=5000)
    ap.add_argument("--ess-chains", type=int, default=64, help="chains used for the ESS estimate of the full run")
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


def cpu_baseline(a, conf, model, niter, warm=1):
    """The oracle (CPU restatement of compiled NIMBLE's sampling path) on the host cores: one chain per thread,
    the one-process-per-core pattern of the NIMBLE manual (chapter_MCMC.Rmd:1223-1229)."""
    import nimble_b200 as nb
    from nimble_b200 import configs
    from oracle.oracle_capi import oracle_backend
    ob = oracle_backend()
    cores = os.cpu_count() or 1
    chains = a.cpu_chains or min(cores, 64)
    cm = nb.compileNimble(nb.buildMCMC(conf), backend=ob)
    dm = cm._ensure(chains)
    inits = np.stack([model.init_vector(i) for i in configs.logistic_inits(a.p, chains)])
    dm.set_values(inits)
    evals_iter = configs.evals_per_iteration(cm)

    def run(n_it, reset):
        ob.check(ob.prepare(dm.h, n_it, 0, 1, 1 if reset else 0))
        ths = [threading.Thread(target=ob.run_range, args=(dm.h, n_it, 0, 1, 1 if reset else 0, 12345, c, c + 1))
               for c in range(chains)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        ob.check(ob.finish(dm.h, n_it, 0, 1))
        return dt
    if warm:
        run(warm, True)
    dt = run(niter, False if warm else True)
    samples = dm.samples()
    val = evals_iter * chains * niter / dt
    return {"value": val, "unit": UNIT, "cores": min(chains, cores), "kind": "port",
            "sample": "%d chains x %d iterations of the same N=%d p=%d workload, one chain per host thread (%d cores)"
                      % (chains, niter, a.n, a.p, cores),
            "ms_per_iteration_per_chain": dt / niter * 1e3, "seconds": dt}, samples, dt, chains


def _emit(line):
    """The JSON line goes to the process's ORIGINAL stdout; everything else that lands on fd 1 meanwhile (NCCL's version
    banner, library chatter) has been diverted to stderr, so stdout carries exactly one line."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if a.impl == "reference":
        if rank != 0:
            return 0
        model, conf, _ = build_model(a)
        a.cpu_iters = max(1, a.steps)
        cb, samples, dt, chains = cpu_baseline(a, conf, model, a.steps, warm=max(1, min(a.warmup, 3)))
        from nimble_b200.ess import min_ess
        mess = min_ess(samples[:, -a.steps:, :])[0] if a.steps >= 20 else None
        line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(a, {"chains": chains, "note": "reference's compiled NIMBLE cannot be built "
                                              "here (no R / Rmath); this is its CPU restatement (oracle/), one chain per core"}),
                "cpu_baseline": cb, "min_ess_per_sec": (mess / dt if mess is not None else None),
                "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        _emit(line)
        return 0

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the B200 back end has no CPU fallback"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    import nimble_b200 as nb
    from nimble_b200 import _capi, configs
    from nimble_b200.ess import min_ess

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
    samples = dm.samples()[:, -a.steps:, :]
    value = evals_iter * C * world * a.steps / (ms_max * 1e-3)

    # ---- end to end through the public API with host buffers: compile (lowering) once, then ONE runMCMC call that
    # uploads the table (X, y, ...) and the inits, runs W+K iterations and reads mvSamples back.
    cm2 = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=first_chain)
    barrier()
    t0 = time.perf_counter()
    out = nb.runMCMC(cm2, niter=a.warmup + a.steps, nburnin=a.warmup, nchains=C, inits=inits_all[first_chain:first_chain + C],
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

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel (fused dbern-logit-inprod block sum), timed alone
        ydecl = len(cm.builder.decls) - 1
        k_ms = dm.time_calculate(ydecl, 10)
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
        # FP64 work of the link + Bernoulli density per evaluation as the kernel implements it (DESIGN.md section 3,
        # ncu profiles/r1f_summary.md): 24.8 FP64-pipe instructions on the reciprocal-free path every realistic linear
        # predictor takes (table-driven exp 11, 1 + e, table-driven log 10, glue 2.8; 35.5 on the general path);
        # one instruction = one multiplier slot = 2 flop in the units of the FMA peak.  Implementation-counted,
        # reported next to the strict figure.
        epi_flops = 2.0 * 24.8 * a.n * C
        ach = contraction_flops / (k_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "k_fused_bern_logit", "achieved": ach, "peak": dmma, "unit": "TFLOP/s",
                "frac": ach / dmma if dmma else None,
                "peak_source": "FP64 mma.sync loop measured live (nimb200_measure_fp64_peaks); MEASURED_PEAKS.json has no FP64 "
                               "figure. DFMA loop: %.1f TFLOP/s; DMMA+DFMA issued together: %.1f TFLOP/s (the two share the FP64 "
                               "multipliers)" % (dfma, mixed),
                "what": "strict algorithmic flops only: the 2*N*p*C contraction",
                # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this kernel at the
                # default workload (profiles/r1f_summary.md); not measurable live, so only quoted for that workload
                "traffic": 429.9e6 if (a.n, a.p, C) == (1_000_000, 50, 1024) else None, "traffic_source": "profiles/r1f_summary.md",
                "algorithmic_bytes": alg_bytes, "kernel_ms": k_ms, "evals_per_s_kernel": a.n * C / (k_ms * 1e-3),
                "fp64_units": {"achieved": (contraction_flops + epi_flops) / (k_ms * 1e-3) / 1e12, "peak": mixed, "unit": "TFLOP/s",
                               "frac": (contraction_flops + epi_flops) / (k_ms * 1e-3) / 1e12 / mixed if mixed else None,
                               "what": "contraction + 24.8 FP64 instructions per evaluation of link and density (as implemented, ncu-counted), "
                                       "against the measured DMMA+DFMA mixed peak"},
                "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                "kernel_share_of_step": k_ms / (ms_max / a.steps)}
        mess, per_col = min_ess(samples)
        cb = None
        if world == 1 and not a.no_cpu_baseline:
            cb, _, _, _ = cpu_baseline(a, conf, model, a.cpu_iters)
        full = None
        if a.full:
            # SURVEY 8d protocol: niter 5000, nburnin 2500, all chains; ESS (coda estimator) on a subset of chains, scaled
            cm3 = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=first_chain)
            d3 = cm3._ensure(C)
            d3.set_values(inits)
            nb_it = a.full_niter // 2
            tfull0 = time.perf_counter()
            d3.run(a.full_niter, nb_it, 1, True, 777)
            torch.cuda.synchronize()
            wall = time.perf_counter() - tfull0
            ms_full, _ = d3.last_timing()
            sf = d3.samples()
            nes = min(a.ess_chains, C)
            mess_f, tot_f = min_ess(sf[:nes])
            scale_c = C / nes
            pooled = sf.reshape(-1, sf.shape[-1])
            cpu_it_per_s = None
            if cb is not None:
                cpu_it_per_s = 1e3 / cb["ms_per_iteration_per_chain"] * cb["cores"]   # chain-iterations per second, all cores
            ess_per_chain_iter = mess_f / nes / a.full_niter
            full = {"niter": a.full_niter, "nburnin": nb_it, "chains": C, "device_seconds": ms_full * 1e-3, "wall_seconds": wall,
                    "min_ess_total": mess_f * scale_c, "min_ess_per_sec": mess_f * scale_c / wall * world,
                    "ess_estimated_on_chains": nes,
                    "cpu_min_ess_per_sec_scaled": (ess_per_chain_iter * cpu_it_per_s if cpu_it_per_s else None),
                    "cpu_scaling_note": "CPU figure = (min ESS per chain-iteration measured on the GPU chains; same algorithm and "
                                        "draws) x (CPU chain-iterations/s of cpu_baseline, all cores): chains are independent",
                    "posterior_mean_first_cols": dict(zip(cm3.colnames[:4], [float(x) for x in pooled.mean(0)[:4]])),
                    "posterior_sd_first_cols": dict(zip(cm3.colnames[:4], [float(x) for x in pooled.std(0)[:4]]))}
            d3.close()
        secondary = None
        if world == 1 and not a.no_secondary:
            # the one HBM-bound logProb kernel of the path, measured live on its own configuration (configs[2]: LMM, N = 1e7,
            # 8 chains): block sum of the data declaration = streaming kernel + reduce/finish launch, CUDA events, min of 3 x 50
            try:
                m3, conf3 = configs.lmm()
                cm_l = nb.compileNimble(nb.buildMCMC(conf3), backend=backend, device=local)
                dl = cm_l._ensure(8)
                dl.set_values(np.tile(m3.init_vector(), (8, 1)))
                yd = len(cm_l.builder.decls) - 1
                nn = cm_l.builder.decls[yd].n
                ms_l = min(dl.time_calculate(yd, 50) for _ in range(3))
                bytes_l = 12.0 * nn
                secondary = [{"kernel": "k_norm_stream_tma + k_norm_reduce_finish", "bound": "hbm",
                              "workload": "configs[2]: y[i] ~ dnorm(mu + u[grp[i]], sd) block sum, N = %d, 8 chains" % nn,
                              "achieved": bytes_l / (ms_l * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                              "frac": bytes_l / (ms_l * 1e-3) / 1e9 / hbm_peak, "us_per_block_sum": ms_l * 1e3,
                              "algorithmic_bytes": bytes_l,
                              "note": "both launches of one block sum are inside the timed region; the streaming kernel alone is "
                                      "23.9-26.3 us under ncu (profiles/r1e_summary.md), 0.70-0.77 of the measured peak"}]
                dl.close()
            except Exception as ex:  # the headline line must not depend on it
                secondary = [{"error": str(ex)[:200]}]
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": workload_config(a, {"chains_total": C * world}),
                "clocks": clk, "gpu_launches": int(launches),
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d / nsteps_e2e, "d2h_bytes_per_step": d2h / nsteps_e2e,
                        "seconds": t_e2e, "steps": nsteps_e2e,
                        "what": "one runMCMC() call on host buffers: table upload + inits + iterations + mvSamples read-back"},
                "roofline": roof, "cpu_baseline": cb,
                "min_ess_per_sec": mess * world / (ms_max * 1e-3),
                "min_ess_note": "sum over chains of coda-style ESS of the timed iterations' samples (adaptation still running); rank 0's chains x n_gpus",
                "evals_per_iteration_per_chain": evals_iter, "host_lowering_seconds": t_build, "full_protocol": full,
                "secondary_rooflines": secondary}
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

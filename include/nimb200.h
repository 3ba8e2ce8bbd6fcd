/*
 * nimb200.h -- C ABI of the B200 multi-chain MCMC back end for NIMBLE models.
 *
 * Everything that crosses this boundary is a plain pointer, a size or a
 * scalar: no torch types, no SEXPs, no C++ classes.  The R glue (.Call shim,
 * see INTEGRATION.md) and the Python host layer (nimble_b200/_capi.py, ctypes)
 * both bind exactly these symbols.
 *
 * Paths below are relative to the reference tree, P/ = packages/nimble/.
 *
 * The unit that crosses is the *node table*: a flat description of a lowered
 * model (what compileNimble turns declInfo / maps / samplerConfs into).  It
 * replaces, for the sampling path only, the generated C++ classes
 *   - nodeFun / indexedNodeInfo            P/inst/include/nimble/nodeFun.h:134-249
 *   - NodeVectorClassNew / NodeInstruction P/inst/include/nimble/accessorClasses.h:50-62
 *   - copierVectorClass / nimCopy          P/inst/CppCode/accessorClasses.cpp:549-572
 *   - the generated sampler classes        P/R/MCMC_samplers.R, P/R/MCMC_conjugacy.R
 */
#ifndef NIMB200_H
#define NIMB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NIMB200_TABLE_VERSION 1
#define NIMB200_MAX_PARAM 3
#define NIMB200_MAX_BLOCK_D 64 /* largest RW_block dimension */

/* ---- distributions: Rmath canonical parameterisation, the "Rdist" column of
 * P/R/distributions_inputList.R:20-149.  The lowering layer appends the
 * reparameterisation ops (sd = 1/sqrt(tau), scale = 1/rate ...) to the
 * parameter programs, exactly where NIMBLE inserts lifted nodes. */
enum {
  NIMB_DNORM = 1,  /* (mean, sd)        inputList.R:144-149 */
  NIMB_DGAMMA = 2, /* (shape, scale)    inputList.R:103-110 */
  NIMB_DPOIS = 3,  /* (lambda)          inputList.R:59-62   */
  NIMB_DBINOM = 4, /* (size, prob); dbern = size 1, inputList.R:20-31 */
  NIMB_DBETA = 5,  /* (shape1, shape2)  inputList.R:70-75   */
  NIMB_DLOGIS = 6, /* (location, scale) inputList.R:139-142 */
  NIMB_DCAT = 7,   /* (prob[1:K]); K in fam_arg[0]; P/inst/CppCode/dists.cpp:914-934 */
  NIMB_DEXP = 8,   /* (rate): dexp_nimble, P/inst/CppCode/dists.cpp:1804-1807; inputList.R:96-100 */
  NIMB_DLNORM = 9  /* (meanlog, sdlog): Rmath dlnorm, sdlog = 1/sqrt(taulog), inputList.R:131-137 */
};

/* ---- expression programs: postfix, evaluated per (node instance i, chain c).
 * One op = one IEEE-754 double operation or one libm call, so the CPU oracle
 * and the device evaluate the same rounding sequence.  Deterministic nodes of
 * the BUGS model (lifted nodes, links: P/R/BUGS_mathCompatibility.R:3-12) are
 * inlined here instead of being stored per chain. */
typedef struct nimb200_op {
  int32_t op;
  int32_t a, b, c;
  double k;
} nimb200_op;

enum {
  NIMB_OP_CONST = 1, /* push k */
  NIMB_OP_DATA = 2,  /* push darr[a][IDX]          (b = index mode, c = its argument) */
  NIMB_OP_CVAR = 3,  /* push state[a]              per-chain scalar */
  NIMB_OP_CVARN = 4, /* push state[a + IDX]        (b = index mode, c = its argument) */
  NIMB_OP_INPROD = 5, /* push sum_{j<k} darr[a][i*c + j] * state[b + j], left to right
                         (inprod: BUGS_mathCompatibility.R:38-39, Utils.h:251) */
  NIMB_OP_DATAJ = 6, /* push darr[a][IDX*c... ] vector element j: darr[a][i*c + j] (dcat prob rows) */
  NIMB_OP_CVARJ = 7, /* push state[a + j]          vector element j (dcat prob in chain state) */
  /* binary */
  NIMB_OP_ADD = 16,
  NIMB_OP_SUB = 17,
  NIMB_OP_MUL = 18,
  NIMB_OP_DIV = 19,
  NIMB_OP_POW = 20,
  /* unary */
  NIMB_OP_NEG = 32,
  NIMB_OP_EXP = 33,
  NIMB_OP_LOG = 34,
  NIMB_OP_SQRT = 35,
  NIMB_OP_ILOGIT = 36, /* 1/(1+exp(-x))  Utils.h:187 */
  NIMB_OP_LOGIT = 37,  /* log(x/(1-x))   Utils.h:191 */
  NIMB_OP_INV = 38,    /* 1/x */
  NIMB_OP_SQ = 39,     /* x*x */
  NIMB_OP_ABS = 40,
  NIMB_OP_LOG1P = 41,
  NIMB_OP_ICLOGLOG = 42, /* 1-exp(-exp(x)) Utils.h:189 */
  NIMB_OP_STEP = 43,     /* x>=0 ? 1 : 0   Utils.h:199 */
  /* dynamic indexing (mu[z[i]]): pop the 1-based index v; v an integer in [1, b] -> push the element, else NaN */
  NIMB_OP_CVAR_IDX = 48, /* push state[a + v - 1] */
  NIMB_OP_DATA_IDX = 49  /* push darr[a][v - 1] */
};

/* index modes for NIMB_OP_DATA / NIMB_OP_CVARN: node index i -> element index */
enum {
  NIMB_IDX_I = 0,    /* i */
  NIMB_IDX_DIV = 1,  /* i / c      (outer index of a nested loop, inner extent c) */
  NIMB_IDX_MOD = 2,  /* i % c */
  NIMB_IDX_IARR = 3  /* iarr[c][i] (0-based group index, e.g. grp[i]) */
};

/* ---- one BUGS declaration = a family of n stochastic node instances
 * (one nodeFun subclass per declaration in the reference,
 * P/R/nimbleFunction_nodeFunctionNew.R:179-216). */
typedef struct nimb200_decl {
  int32_t dist;       /* NIMB_D* */
  int32_t n;          /* node instances; node i is instance i of the declaration */
  int32_t value_kind; /* 0: data, value = darr[value_ref][i]; 1: per-chain, state[value_ref + i] */
  int32_t value_ref;
  int32_t nparam;
  int32_t prog_off[NIMB200_MAX_PARAM]; /* parameter programs, offsets into ops[] */
  int32_t prog_len[NIMB200_MAX_PARAM];
  int32_t family;     /* NIMB_FAM_*: 0 = generic interpreter, else a fused kernel the host recognised */
  int32_t fam_arg[6];
} nimb200_decl;

enum {
  NIMB_FAM_GENERIC = 0,
  /* y[i] ~ dbern(ilogit(inprod(X[i,1:p], beta[1:p]))):
   * fam_arg = {X darr id, beta slot base, p, row stride of X} */
  NIMB_FAM_BERN_LOGIT_INPROD = 1,
  /* y[i] ~ dpois(lam[G(i)] * exp(b0 + b1 * x[i])), G an index map (identity / i div c / int array):
   * fam_arg = {x darr id, lam slot base, index mode, index argument, b0 slot, b1 slot} */
  NIMB_FAM_POIS_LOGLIN = 2,
  /* y[i] ~ dnorm(A + v[G(i)], sd = S), A and S chain-level programs (no node index), v gathered through an index map:
   * fam_arg = {ops offset of A, length of A, v slot base, index mode (-1: no gathered term), index argument, 0};
   * S is the declaration's own second parameter program */
  NIMB_FAM_NORM_GATHER = 3,
  /* v[i] ~ dbern(gate[G(i)] * ilogit(c0 + c1 * x[i])), v data or latent, gate an optional latent 0/1 vector:
   * fam_arg = {x darr id, c0 slot, c1 slot, gate slot base (-1: none), gate index mode, gate index argument} */
  NIMB_FAM_BERN_LOGIT_LIN = 4
};

/* ---- node selections: which instances of a declaration depend on target t */
enum {
  NIMB_SEL_ALL = 0,    /* every instance, for every target */
  NIMB_SEL_SAME = 1,   /* instance a + t */
  NIMB_SEL_STRIDE = 2, /* instances t*a .. t*a + a - 1 */
  NIMB_SEL_SEG = 3,    /* instances iarr[a][t] .. iarr[a][t+1]-1 (sorted groups) */
  NIMB_SEL_CSR = 4,    /* instances iarr[b][ iarr[a][t] .. iarr[a][t+1]-1 ] */
  NIMB_SEL_RANGE = 5   /* instances a .. b-1, for every target */
};

/* parameter codes for nimb200_dep.param: 0..2 = canonical parameter index; the alternative
 * parameters NIMBLE derives from the canonical ones (altParams, distributions_inputList.R:106,147) */
enum {
  NIMB_PARAM_RATE = 16, /* 1/scale   of dgamma(shape, scale) */
  NIMB_PARAM_TAU = 17   /* sd^-2     of dnorm(mean, sd) */
};

/* conjugate link of a dependent on the target, P/R/MCMC_conjugacy.R:1397-1524 */
enum { NIMB_LINK_IDENTITY = 0, NIMB_LINK_ADDITIVE = 1, NIMB_LINK_MULTIPLICATIVE = 2, NIMB_LINK_LINEAR = 3 };

typedef struct nimb200_dep {
  int32_t decl;
  int32_t sel, a, b;
  int32_t stage; /* 0: the target's own prior node(s) ("ProposalStage", MCMC_samplers.R:660-663);
                    1: stochastic dependents ("DepStage") */
  int32_t link;  /* conjugate samplers only */
  int32_t param; /* conjugate samplers only: canonical parameter index of the dependent that the
                    target enters (lambda, mean, scale<-rate ...), probed as in MCMC_conjugacy.R:766-820 */
  int32_t pad;
} nimb200_dep;

/* ---- samplers: one entry = n_targets conditionally independent scalar targets that the
 * reference would visit one after another with the same sampler type (or one RW_block). */
enum {
  NIMB_SAMPLER_RW = 1,         /* MCMC_samplers.R:320-461 */
  NIMB_SAMPLER_RW_BLOCK = 2,   /* MCMC_samplers.R:643-810 */
  NIMB_SAMPLER_BINARY = 3,     /* MCMC_samplers.R:115-152 */
  NIMB_SAMPLER_CONJ_GAMMA = 4, /* dgamma prior; dpois / dgamma(rate) / dnorm(tau) dependents, MCMC_conjugacy.R:80-93 */
  NIMB_SAMPLER_CONJ_NORMAL = 5, /* dnorm prior; dnorm(mean) dependents, MCMC_conjugacy.R:109-115 */
  NIMB_SAMPLER_CONJ_BETA = 6,   /* dbeta prior; dbern / dbin(prob) dependents, identity link, MCMC_conjugacy.R:4-15 */
  NIMB_SAMPLER_POSTERIOR_PREDICTIVE = 7, /* MCMC_samplers.R:86-105: simulate the target from its own distribution (dnorm, dgamma,
                                            dbeta, dexp, dlogis, dbin, dpois with mean <= 500).  The reference's sampler simulates the
                                            whole predictive network downstream of its target in topological order; the host emits one
                                            table sampler per member declaration, parents first, so a sequence of these is that
                                            sampler.  Stage-1 entries name the dependents whose cached block sums the draw invalidates */
  NIMB_SAMPLER_CATEGORICAL = 9, /* MCMC_samplers.R:162-226: dcat target with K <= NIMB200_MAX_CATEGORIES categories: the conditional
                                   log-density of every category, then rcat (dists.cpp:936-961) on exp(logProb - max) */
  NIMB_SAMPLER_SLICE = 8 /* MCMC_samplers.R:864-972: stepping-out / shrinkage slice sampler with width adaptation; discrete
                            targets (dpois, dbin, dcat) are floored.  scale = sliceWidth, tries = sliceMaxSteps,
                            pad = maxContractions, adaptive / adapt_interval as for RW */
};

typedef struct nimb200_sampler {
  int32_t type;
  int32_t n_targets;   /* scalar samplers: T targets at slots target_slot + t; RW_block: 1 */
  int32_t target_slot;
  int32_t d;           /* RW_block dimension (target slots target_slot .. +d-1); else 1 */
  int32_t dep_off, n_dep; /* into deps[] (stage-0 entries first) */
  int32_t log_scale;   /* RW: control$log */
  int32_t reflective;  /* RW: control$reflective */
  int32_t adaptive;
  int32_t adapt_interval;
  int32_t adapt_scale_only; /* RW_block */
  int32_t tries;            /* RW_block */
  int32_t propcov_darr;     /* RW_block: darr id of a d*d column-major matrix, -1 = identity */
  int32_t pad;
  double scale;
  double adapt_factor_exponent;
  double lower, upper;      /* reflective RW bounds */
} nimb200_sampler;

typedef struct nimb200_table {
  int32_t version;  /* NIMB200_TABLE_VERSION */
  int32_t n_slots;  /* doubles of per-chain state (all non-data stochastic nodes) */
  int32_t n_darr;   /* shared (chain-invariant) double arrays: data and constants */
  int32_t n_iarr;   /* shared int32 arrays: group indices, segment offsets */
  const double *const *darr;
  const int64_t *darr_len;
  const int32_t *const *iarr;
  const int64_t *iarr_len;
  int32_t n_ops;
  int32_t n_decl;
  const nimb200_op *ops;
  const nimb200_decl *decl;
  int32_t n_samplers;
  int32_t n_deps;
  const nimb200_sampler *samplers; /* in samplerExecutionOrder, MCMC_build.R:287-290 */
  const nimb200_dep *deps;
  int32_t n_monitors;
  int32_t pad;
  const int32_t *monitor_slots; /* state slots copied to mvSamples, MCMC_build.R:294-303 */
} nimb200_table;

typedef struct nimb200_model nimb200_model; /* opaque; owns all device buffers */

/* Every function returns 0 on success, non-zero on error; the message is read with
 * nimb200_last_error() (thread-local).  Replaces NIMERROR/Rf_error, Utils.h:60. */
const char *nimb200_last_error(void);
int nimb200_version(void);

/* compileNimble(model, mcmc) for the sampling path: validates the table, uploads the shared
 * arrays once, allocates chain-major state for nchains chains on CUDA device `device`.
 * Replaces populateNodeFxnVectorNew_internal / populateIndexedNodeInfoTable / populateCopierVector,
 * P/inst/CppCode/accessorClasses.cpp:961-1060.  first_chain is the global index of chain 0 of this
 * model (chain-sharded multi-GPU runs); it only enters the RNG key. */
int nimb200_model_create(const nimb200_table *table, int32_t nchains, int32_t first_chain, int32_t device,
                         nimb200_model **out);
int nimb200_model_free(nimb200_model *m);

/* model$setInits / values(model, nodes) <- v for every chain at once,
 * P/inst/CppCode/accessorClasses.cpp:513-545.  Host layout: [nchains][n_slots]. */
int nimb200_set_values(nimb200_model *m, const double *values);
int nimb200_get_values(nimb200_model *m, double *values);

/* model$calculate(nodes): sum of log-densities of instances [lo,hi) of one declaration, one value per
 * chain (P/inst/CppCode/accessorClasses.cpp:49-57).  When the range is the whole declaration the
 * per-chain block sum is cached as that declaration's logProb. */
int nimb200_calculate(nimb200_model *m, int32_t decl, int32_t lo, int32_t hi, double *out_nchains);
/* same, but one value per node instance and chain: out[(i-lo)*nchains + c] (the logProb_x[i] the
 * reference stores, nimbleFunction_nodeFunctionNew.R:207). */
int nimb200_calculate_nodes(nimb200_model *m, int32_t decl, int32_t lo, int32_t hi, double *out);
/* model$calculateDiff(nodes) on a whole declaration: new block sum minus the cached one, cache updated
 * (nodeFunctionNew.R:208-209, accessorClasses.cpp:237-245). */
int nimb200_calculateDiff(nimb200_model *m, int32_t decl, double *out_nchains);
/* model$getLogProb(nodes) on a whole declaration: the cached block sum (accessorClasses.cpp:256-264). */
int nimb200_getLogProb(nimb200_model *m, int32_t decl, double *out_nchains);

/* Cmcmc$run(niter, nburnin, thin, reset): MCMC_build.R:217-321.  All chains advance together.
 * reset != 0: samplers reset, iteration counter and sample buffer cleared, the model is initialised
 * (every node calculated once, initializeModel.R:78-96).  reset == 0 continues the chains and appends.
 * seed keys the counter-based RNG (DESIGN.md "draws"). */
int nimb200_mcmc_run(nimb200_model *m, int64_t niter, int64_t nburnin, int32_t thin, int32_t reset,
                     uint64_t seed);
/* as.matrix(mvSamples) for every chain: out[c][row][col], rows = saved samples, cols = monitors
 * (types_util.R:199-214).  *nrows_out receives the row count; out may be NULL to query it. */
int nimb200_mcmc_samples(nimb200_model *m, double *out, int64_t *nrows_out);
/* nimbleOptions(MCMCsaveHistory = TRUE) (P/R/MCMC_samplers.R:394-399, 726-739; getScaleHistory / getAcceptanceHistory /
 * getPropCovHistory :421-448, 775-780): once enabled, every adaptation of every RW / RW_block sampler records the scale it used,
 * the interval's acceptance rate and -- RW_block with d <= 10 (maxDimCovHistory) -- the proposal covariance it used, for up to
 * max_adaptations adaptations since the last reset.  history(): what = 0 scaleHistory, 1 acceptanceHistory -> out[adaptation][T][C];
 * 2 propCovHistory -> out[adaptation][d*d][C]; *n_adapt_out = adaptations recorded (out may be NULL to query it). */
int nimb200_sampler_history_enable(nimb200_model *m, int32_t max_adaptations);
int nimb200_sampler_history(nimb200_model *m, int32_t sampler, int32_t what, double *out, int64_t cap, int32_t *n_adapt_out);
/* samplesSummary(samples) (P/R/MCMC_utils.R:315-341) computed on the device, without reading mvSamples back: out[5][n_monitors] =
 * Mean, Median, St.Dev., 95%CI_low, 95%CI_upp of every monitored column over all saved rows of the chains [chain_lo, chain_hi)
 * pooled (one chain: that chain's summary; all chains: runMCMC's `all.chains`).  Quantiles are exact order statistics with R's
 * default (type 7) interpolation. */
int nimb200_samples_summary(nimb200_model *m, int32_t chain_lo, int32_t chain_hi, double *out);
/* device-side time of the last nimb200_mcmc_run in milliseconds (CUDA events on the run stream),
 * and how many kernels it launched. */
int nimb200_mcmc_last_timing(nimb200_model *m, double *ms_out, int64_t *launches_out);
/* sampler state read-back (scale, acceptance counters, proposal covariance): the fields the reference
 * keeps in the sampler objects, MCMC_samplers.R:449-459, 793-808.
 * what: 0 scale [T][C]; 1 timesAdapted [T][C]; 2 propCov [d*d][C] (RW_block). */
int nimb200_sampler_state(nimb200_model *m, int32_t sampler, int32_t what, double *out, int64_t cap);

/* Data-node sharding across the GPUs of one node (the reference has no counterpart, one core walks all
 * N nodes).  Every rank creates the same model (same table, chains and seed), then calls nimb200_enable_data_sharding:
 * data declarations with >= min_nodes instances are split into contiguous slices, and each block sum / conjugate
 * contribution that involves a sharded declaration is all-reduced (sum, FP64) over NCCL inside the library, after
 * which all ranks take identical decisions.  id128 is the 128-byte ncclUniqueId produced by nimb200_nccl_unique_id on
 * one rank and distributed by the host (NULL: slices without a communicator, block sums stay per-rank partial sums). */
int nimb200_nccl_unique_id(void *id128);
int nimb200_enable_data_sharding(nimb200_model *m, int32_t nranks, int32_t rank, const void *id128, int32_t min_nodes);

/* The same data-node sharding with the all-reduce done by a kernel of this library over NVLink peer memory instead of
 * NCCL (no reference counterpart either: MCMC_run.R:127-153 is one sequential process).  The per-proposal message is
 * `chains` doubles, so what matters is latency: every rank pushes its partial sums into every rank's exchange buffer with
 * plain stores, publishes an epoch flag, waits for its peers' flags in its own memory and adds the contributions in rank
 * order (identical bits, hence identical decisions, on all ranks).  Being ordinary kernels, these all-reduces are captured
 * into the iteration's CUDA graph with everything else.  Protocol, 2 to 8 ranks of ONE node, one model per rank (separate
 * processes, or several devices of one process):
 *   1. every rank: nimb200_peer_export(m, nranks, handle) -- allocates and clears this rank's exchange buffer and writes an
 *      opaque NIMB200_PEER_HANDLE_BYTES-byte handle;
 *   2. the host gathers all handles to all ranks, in rank order (any all-gather; it doubles as the barrier that guarantees
 *      every buffer is cleared before any peer writes to it);
 *   3. every rank: nimb200_enable_data_sharding_peer(m, nranks, rank, handles[nranks][128], min_nodes).
 * A rank whose peers do not arrive within a few seconds gives up (the run then returns an error) instead of spinning. */
#define NIMB200_PEER_HANDLE_BYTES 128
int nimb200_peer_export(nimb200_model *m, int32_t nranks, void *handle128);
int nimb200_enable_data_sharding_peer(nimb200_model *m, int32_t nranks, int32_t rank, const void *handles, int32_t min_nodes);
/* all-reduces enqueued by the last nimb200_mcmc_run (either transport), and whether its iterations were replayed from a
 * captured CUDA graph; measurement only. */
int nimb200_mcmc_last_allreduces(nimb200_model *m, int64_t *count_out, int32_t *graph_replayed_out);
/* average device milliseconds of one all-reduce of `count` doubles through the data-sharded model's transport (every rank of
 * the group makes the same call); measurement only. */
int nimb200_time_allreduce(nimb200_model *m, int32_t count, int32_t reps, double *ms_avg);

/* measurement hooks for bench.py (no counterpart in the reference, whose only profiling hook is
 * run(time=TRUE), MCMC_build.R:281-285): average device milliseconds of one launch of the block-sum kernel
 * behind nimb200_calculate(decl, whole range), and live FP64 peaks (DFMA loop, FP64 mma.sync loop). */
int nimb200_time_calculate(nimb200_model *m, int32_t decl, int32_t reps, double *ms_avg);
int nimb200_measure_fp64_peaks(int device, double *dfma_tflops, double *dmma_tflops);
/* same, for a loop that issues FP64 mma.sync and DFMA together (equal flops of each): tells whether the FP64 tensor
 * sub-pipe and the FP64 pipe add up or share their multipliers. */
int nimb200_measure_fp64_mixed(int device, double *tflops);
/* DFMA TFLOP/s of a loop with 0, 4, 8 and 16 independent integer instructions per 8 DFMAs (tflops4[4]): how much of the FP64
 * pipe a mixed instruction stream can use -- the issue model behind the family kernels' instruction budgets. */
int nimb200_measure_fp64_issue_mix(int device, double *tflops4);
/* diagnostic: the fused kernels' own exp (which=0) / log (1) / reciprocal (2) / dbern-logit log-density for y=0 (3),
 * y=1 (4) applied to a host array, so that tests can measure their accuracy against libm. */
int nimb200_debug_math(int32_t which, const double *x, int32_t n, double *out);
/* diagnostic: the tcgen05 experiment (csrc/ozaki_logit.cu; no reference counterpart): the dbern-logit-inprod block sum with the
 * contraction X.beta computed as exact int8 slice products on the 5th-generation tensor cores (tcgen05.mma kind::i8,
 * accumulators in TMEM) and recombined in FP64.  HOST arrays X [n][p] row-major, y [n], beta [p][C] (chain fastest); eta_out
 * [n][C], sums_out [C] and ms_out (average milliseconds of one launch) may each be NULL.  Allocates and frees its own device
 * buffers.  Not used by the sampling path (measured slower than the FP64 mma.sync kernel). */
int nimb200_debug_ozaki(const double *X, const double *y, int32_t n, int32_t p, const double *beta, int32_t C, double *eta_out,
                        double *sums_out, double *ms_out);
/* diagnostic: clock64 timeline of CTA (0, 0) of that kernel.  First call (out = NULL) switches the stamps on; after one or more
 * nimb200_debug_ozaki calls, the second call copies them out ([18 warps][64 tiles][2 halves][4 events], int64) and switches
 * them off again (scripts/probe_ozaki_trace.py). */
int nimb200_debug_ozaki_trace(long long *out_or_null);

/* No counterpart in the reference (NIMBLE has no device code and nothing in it is asked to run without
 * allocating).  Per-chain block log-probability of y[i] ~ dbern(ilogit(inprod(X[i,1:p], beta[1:p]))) -- the quantity
 * nimb200_calculate() returns for such a declaration -- for callers that must not allocate device memory: a host
 * program whose own memory manager owns the whole GPU (start-up arena, fixed per-worker budget), or a steady-state
 * loop that cannot afford the device-wide synchronisation of cudaMalloc/cudaFree.  Needs no nimb200_model.
 * All pointers are DEVICE pointers supplied by the caller: X row-major [n][ld] (first p columns used), y [n] in {0,1},
 * beta [p][nchains] (chain index fastest, like the model state), out [nchains].  The work is enqueued on cuda_stream
 * (a cudaStream_t, NULL = default stream) and is NOT synchronised.  The kernel's scratch is a built-in device array:
 * nchains <= NIMB200_NOALLOC_MAX_CHAINS and p <= 64, otherwise an error is returned; the array (2.3 MiB) stays
 * resident on every GPU that loaded the library; the first call on a device looks its address up, so make that call
 * before starting any stream capture; calls share the array, so calls that overlap on one GPU race (each call
 * rewrites all of the array it reads, nothing needs clearing in between).  Calls no cudaMalloc/cudaFree. */
#define NIMB200_MAX_CATEGORIES 64 /* categorical sampler: categories per target */
#define NIMB200_NOALLOC_MAX_CHAINS 2048
int nimb200_bern_logit_block_logprob_noalloc(const double *d_X, int64_t ld, const double *d_y, int32_t n, int32_t p,
                                             const double *d_beta, int32_t nchains, double *d_out, void *cuda_stream);

/* run(time = TRUE) / getTimes() (P/R/MCMC_build.R:281-285, 323-326): with timing enabled every sampler's launches are bracketed by
 * CUDA events on the run stream and the elapsed times accumulated per sampler (milliseconds, all chains together; the run is then
 * not captured into a CUDA graph and synchronises once per iteration, so it is a profiling mode).  sampler_times fills
 * out_ms[n_samplers] with the totals since timing was last enabled. */
int nimb200_mcmc_set_timing(nimb200_model *m, int32_t enable);
int nimb200_mcmc_sampler_times(nimb200_model *m, double *out_ms, int32_t n);

/* online conditional WAIC without grouping (P/R/MCMC_WAIC.R:209-282; call site MCMC_build.R:304-310): once enabled, every
 * post-burn-in iteration of every chain updates, per data node, a running log-sum-exp of its log-density (lppd) and a Welford
 * variance (pWAIC).  nimb200_mcmc_run(reset = 1) clears the statistics.  waic_get: out[0] = WAIC = -2 (lppd - pWAIC),
 * out[1] = lppd, out[2] = pWAIC, out[3] = number of (iteration, chain) samples used.  Data-sharded models: nimb200_waic_partial below. */
int nimb200_waic_enable(nimb200_model *m, int32_t enable);
int nimb200_waic_get(nimb200_model *m, double *out4);
/* data-sharded models: every rank accumulates the statistics of its own slice of the data nodes; out4 = {sum over those nodes
 * of log(sum exp lp) (before the -n log(count) term), sum of the per-node sums of squares, number of nodes, count}.  The
 * caller adds the first three over the ranks of the group: lppd = S0 - S2 log(count), pWAIC = S1 / (count - 1). */
int nimb200_waic_partial(nimb200_model *m, double *out4);

/* initializeModel (P/R/initializeModel.R:78-139): every non-data stochastic node whose value is NaN (a missing initial value)
 * is simulated from its own distribution given its parents, declaration by declaration in the order given (parents first);
 * variates come from the counter (seed, chain; iteration 0, sampler id 0x7fff0000 + declaration, target = node).  Supported
 * for dnorm, dgamma, dbeta, dexp, dlogis, dbin and dpois (mean <= 500) nodes; a NaN left in any other declaration is an error. */
int nimb200_initialize_model(nimb200_model *m, const int32_t *decl_order, int32_t n_order, uint64_t seed);

/* draws: fills out[n] with the standard normals (kind 0), uniforms (kind 1), gamma(shape a, scale 1)
 * variates (kind 2), beta(a, 1 + a/2) variates (kind 3), Poisson(a) variates (kind 4) or binomial(25, a) variates (kind 5) that sampler `sampler` would consume on the device for (chain, iter, target), so the
 * host can replay identical draws into another implementation. */
int nimb200_draws(uint64_t seed, int32_t chain, int64_t iter, int32_t sampler, int32_t target,
                  int32_t kind, double a, int32_t n, double *out);

#ifdef __cplusplus
}
#endif
#endif /* NIMB200_H */

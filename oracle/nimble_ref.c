/*
 * nimble_ref.c -- TEST INFRASTRUCTURE (parity oracle and timed CPU baseline).  Not part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Sequential, one-chain-at-a-time CPU restatement of NIMBLE's compiled sampling path, driven by the same
 * node table (include/nimb200.h) the device library consumes.  It keeps the reference's structure:
 *   - one stored logProb per node instance and chain (logProb_x[i], nimbleFunction_nodeFunctionNew.R:207-210)
 *   - calculate / calculateDiff / getLogProb as left-to-right loops over node instances
 *     (P/inst/CppCode/accessorClasses.cpp:49-57, 237-245, 256-264)
 *   - a saved copy of values and logProbs (mvSaved row 1) that every sampler synchronises node by node on
 *     accept and on reject (nimCopy, accessorClasses.cpp:549-572; MCMC_samplers.R:374-383)
 *   - sampler_RW (MCMC_samplers.R:320-461), sampler_RW_block (:643-810, calcAdaptationFactor
 *     MCMC_utils.R:177-205), sampler_binary (:115-152), conjugate dgamma / dnorm samplers with the numeric
 *     offset/coeff probing (MCMC_conjugacy.R:80-93, 109-115, 651-692, 721-1034), decide (Utils.cpp:54-63)
 *   - MCMC::run (MCMC_build.R:217-321) with initializeModel (initializeModel.R:78-139), chains run one after
 *     another as runMCMC does (MCMC_run.R:127-153).
 * Deterministic nodes are inlined into the parameter programs (same arithmetic, same order of operations).
 * Random draws come from draws_ref.c, not from R's RNG (see there).
 * R could not be run in the build container (no R, no Rmath): parity of the scalar densities is pinned to
 * mpmath, parity of the samplers to the known-answer posteriors in the reference's tests (tests/golden/).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../include/nimb200.h"
#include "draws_ref.h"
#include "nmath_ref.h"

typedef struct {
  double scale, scale0, gamma1, sumJumps;
  long timesRan, timesAccepted, timesAdapted;
  double *propCov, *propCov0, *chol, *cholScale, *empirSamp; /* RW_block, column-major d x d / adaptInterval x d */
} SState;

typedef struct ref_model {
  nimb200_table t;
  double **darr;
  int32_t **iarr;
  int C, first_chain;
  double *state, *saved;       /* [C][n_slots] */
  double **lp, **lp_saved;     /* per decl [C][n] */
  SState **ss;                 /* per sampler [T*C] */
  double *samples;             /* [C][cap][n_mon] */
  int64_t nrows, cap;
  int64_t iter_done;
  uint64_t evals;              /* stochastic node evaluations (logProb evals) */
  int waic;                    /* online WAIC enabled (single-threaded runs only) */
  double **waic_acc;           /* per data declaration [4][n]: max, scaled sum, mean, sum of squares (MCMC_WAIC.R:186-193) */
  int64_t waic_count;          /* mcmcIter */
} ref_model;

static __thread uint64_t tl_evals; /* per-thread evaluation counter (flushed into the model at the end of a run) */
static char g_err[512];
const char *ref_last_error(void) { return g_err; }
static int fail(const char *msg) { snprintf(g_err, sizeof g_err, "%s", msg); return 1; }

/* ------------------------------------------------------------------ expressions */
static inline int idx_of(const ref_model *m, int mode, int arg, int i) {
  switch (mode) {
    case NIMB_IDX_I: return i;
    case NIMB_IDX_DIV: return i / arg;
    case NIMB_IDX_MOD: return i % arg;
    default: return m->iarr[arg][i];
  }
}

static double eval_prog(const ref_model *m, const double *st, int off, int len, int i, int j) {
  double stk[32];
  int sp = 0, q;
  const nimb200_op *ops = m->t.ops + off;
  for (q = 0; q < len; q++) {
    const nimb200_op *o = ops + q;
    switch (o->op) {
      case NIMB_OP_CONST: stk[sp++] = o->k; break;
      case NIMB_OP_DATA: stk[sp++] = m->darr[o->a][idx_of(m, o->b, o->c, i)]; break;
      case NIMB_OP_CVAR: stk[sp++] = st[o->a]; break;
      case NIMB_OP_CVARN: stk[sp++] = st[o->a + idx_of(m, o->b, o->c, i)]; break;
      case NIMB_OP_INPROD: {
        const double *x = m->darr[o->a] + (size_t)i * o->c;
        const double *v = st + o->b;
        int p = (int)o->k, jj;
        double s = 0.0;
        for (jj = 0; jj < p; jj++) s += x[jj] * v[jj];
        stk[sp++] = s;
        break;
      }
      case NIMB_OP_DATAJ: stk[sp++] = m->darr[o->a][(size_t)i * o->c + j]; break;
      case NIMB_OP_CVARJ: stk[sp++] = st[o->a + j]; break;
      case NIMB_OP_CVAR_IDX: {
        double v = stk[sp - 1];
        stk[sp - 1] = (v >= 1.0 && v <= (double)o->b && v == floor(v)) ? st[o->a + (int)v - 1] : NAN;
        break;
      }
      case NIMB_OP_DATA_IDX: {
        double v = stk[sp - 1];
        stk[sp - 1] = (v >= 1.0 && v <= (double)o->b && v == floor(v)) ? m->darr[o->a][(int)v - 1] : NAN;
        break;
      }
      case NIMB_OP_ADD: sp--; stk[sp - 1] = stk[sp - 1] + stk[sp]; break;
      case NIMB_OP_SUB: sp--; stk[sp - 1] = stk[sp - 1] - stk[sp]; break;
      case NIMB_OP_MUL: sp--; stk[sp - 1] = stk[sp - 1] * stk[sp]; break;
      case NIMB_OP_DIV: sp--; stk[sp - 1] = stk[sp - 1] / stk[sp]; break;
      case NIMB_OP_POW: sp--; stk[sp - 1] = pow(stk[sp - 1], stk[sp]); break;
      case NIMB_OP_NEG: stk[sp - 1] = -stk[sp - 1]; break;
      case NIMB_OP_EXP: stk[sp - 1] = exp(stk[sp - 1]); break;
      case NIMB_OP_LOG: stk[sp - 1] = log(stk[sp - 1]); break;
      case NIMB_OP_SQRT: stk[sp - 1] = sqrt(stk[sp - 1]); break;
      case NIMB_OP_ILOGIT: stk[sp - 1] = 1. / (1. + exp(-stk[sp - 1])); break;
      case NIMB_OP_LOGIT: stk[sp - 1] = log(stk[sp - 1] / (1. - stk[sp - 1])); break;
      case NIMB_OP_INV: stk[sp - 1] = 1. / stk[sp - 1]; break;
      case NIMB_OP_SQ: stk[sp - 1] = stk[sp - 1] * stk[sp - 1]; break;
      case NIMB_OP_ABS: stk[sp - 1] = fabs(stk[sp - 1]); break;
      case NIMB_OP_LOG1P: stk[sp - 1] = log1p(stk[sp - 1]); break;
      case NIMB_OP_ICLOGLOG: stk[sp - 1] = 1. - exp(-exp(stk[sp - 1])); break;
      case NIMB_OP_STEP: stk[sp - 1] = stk[sp - 1] >= 0 ? 1. : 0.; break;
      default: return NAN;
    }
  }
  return stk[0];
}

static inline double node_value(const ref_model *m, const double *st, const nimb200_decl *d, int i) {
  return d->value_kind ? st[d->value_ref + i] : m->darr[d->value_ref][i];
}

/* getParam_0D_double: accessorClasses.h:210-224, switch generated by nodeFunctionNew.R:379-412 */
static double get_param(const ref_model *m, const double *st, int di, int i, int code) {
  const nimb200_decl *d = m->t.decl + di;
  if (code < NIMB200_MAX_PARAM) return eval_prog(m, st, d->prog_off[code], d->prog_len[code], i, 0);
  if (code == NIMB_PARAM_RATE) return 1.0 / eval_prog(m, st, d->prog_off[1], d->prog_len[1], i, 0);
  if (code == NIMB_PARAM_TAU) return pow(eval_prog(m, st, d->prog_off[1], d->prog_len[1], i, 0), -2.0);
  return NAN;
}

/* one nodeFun::calculate body: d<dist>(value, params, log = 1) */
static double calc_node(ref_model *m, const double *st, int di, int i) {
  const nimb200_decl *d = m->t.decl + di;
  double x = node_value(m, st, d, i), p0 = 0, p1 = 0;
  tl_evals++;
  if (d->dist == NIMB_DCAT) {
    int K = d->fam_arg[0], j;
    double pr[256], *pp = pr, r;
    if (K > 256) pp = (double *)malloc(sizeof(double) * K);
    for (j = 0; j < K; j++) pp[j] = eval_prog(m, st, d->prog_off[0], d->prog_len[0], i, j);
    r = ref_dcat(x, pp, K);
    if (pp != pr) free(pp);
    return r;
  }
  p0 = eval_prog(m, st, d->prog_off[0], d->prog_len[0], i, 0);
  if (d->nparam > 1) p1 = eval_prog(m, st, d->prog_off[1], d->prog_len[1], i, 0);
  switch (d->dist) {
    case NIMB_DNORM: return ref_dnorm(x, p0, p1);
    case NIMB_DGAMMA: return ref_dgamma(x, p0, p1);
    case NIMB_DPOIS: return ref_dpois(x, p0);
    case NIMB_DBINOM: return ref_dbinom(x, p0, p1);
    case NIMB_DBETA: return ref_dbeta(x, p0, p1);
    case NIMB_DLOGIS: return ref_dlogis(x, p0, p1);
    case NIMB_DEXP: return ref_dexp_nimble(x, p0);
    case NIMB_DLNORM: return ref_dlnorm(x, p0, p1);
  }
  return NAN;
}

/* ------------------------------------------------------------------ selections */
typedef struct { int lo, hi; const int32_t *idx; } Sel;
static Sel sel_of(const ref_model *m, const nimb200_dep *dp, int t) {
  Sel s; s.idx = NULL;
  switch (dp->sel) {
    case NIMB_SEL_ALL: s.lo = 0; s.hi = m->t.decl[dp->decl].n; break;
    case NIMB_SEL_SAME: s.lo = dp->a + t; s.hi = s.lo + 1; break;
    case NIMB_SEL_STRIDE: s.lo = t * dp->a; s.hi = s.lo + dp->a; break;
    case NIMB_SEL_SEG: s.lo = m->iarr[dp->a][t]; s.hi = m->iarr[dp->a][t + 1]; break;
    case NIMB_SEL_CSR: s.lo = m->iarr[dp->a][t]; s.hi = m->iarr[dp->a][t + 1]; s.idx = m->iarr[dp->b]; break;
    default: s.lo = dp->a; s.hi = dp->b; break;
  }
  return s;
}

#define FOR_DEP_NODES(m, sp, t, stagesel, ...)                                    \
  do {                                                                            \
    int _q;                                                                       \
    for (_q = 0; _q < (sp)->n_dep; _q++) {                                        \
      const nimb200_dep *dp = (m)->t.deps + (sp)->dep_off + _q;                   \
      if ((stagesel) >= 0 && dp->stage != (stagesel)) continue;                   \
      Sel _s = sel_of((m), dp, (t));                                              \
      int _k;                                                                     \
      for (_k = _s.lo; _k < _s.hi; _k++) {                                        \
        int node = _s.idx ? _s.idx[_k] : _k;                                      \
        int di = dp->decl;                                                        \
        __VA_ARGS__                                                               \
      }                                                                           \
    }                                                                             \
  } while (0)

/* model$calculate(nodes): accessorClasses.cpp:49-57 */
static double calculate_deps(ref_model *m, int c, const nimb200_sampler *sp, int t, int stage) {
  const double *st = m->state + (size_t)c * m->t.n_slots;
  double ans = 0;
  FOR_DEP_NODES(m, sp, t, stage, {
    double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n;
    lp[node] = calc_node(m, st, di, node);
    ans += lp[node];
  });
  return ans;
}
/* model$calculateDiff(nodes): accessorClasses.cpp:237-245, nodeFunctionNew.R:208-209 */
static double calculateDiff_deps(ref_model *m, int c, const nimb200_sampler *sp, int t, int stage) {
  const double *st = m->state + (size_t)c * m->t.n_slots;
  double ans = 0;
  FOR_DEP_NODES(m, sp, t, stage, {
    double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n;
    double nw = calc_node(m, st, di, node);
    ans += nw - lp[node];
    lp[node] = nw;
  });
  return ans;
}
/* model$getLogProb(nodes): accessorClasses.cpp:256-264 */
static double getLogProb_deps(ref_model *m, int c, const nimb200_sampler *sp, int t, int stage) {
  double ans = 0;
  FOR_DEP_NODES(m, sp, t, stage, {
    const double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n;
    ans += lp[node];
  });
  return ans;
}
/* nimCopy of logProbs between model and mvSaved, one node at a time (accessorClasses.cpp:549-572) */
static void copy_lp(ref_model *m, int c, const nimb200_sampler *sp, int t, int stage, int to_saved) {
  FOR_DEP_NODES(m, sp, t, stage, {
    double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n;
    double *ls = m->lp_saved[di] + (size_t)c * m->t.decl[di].n;
    if (to_saved) ls[node] = lp[node]; else lp[node] = ls[node];
  });
}
static void copy_targets(ref_model *m, int c, const nimb200_sampler *sp, int t, int to_saved) {
  double *st = m->state + (size_t)c * m->t.n_slots, *sv = m->saved + (size_t)c * m->t.n_slots;
  int j, d = sp->type == NIMB_SAMPLER_RW_BLOCK ? sp->d : 1, s0 = sp->target_slot + (sp->type == NIMB_SAMPLER_RW_BLOCK ? 0 : t);
  for (j = 0; j < d; j++) { if (to_saved) sv[s0 + j] = st[s0 + j]; else st[s0 + j] = sv[s0 + j]; }
}

/* decide(): P/inst/CppCode/Utils.cpp:54-63 */
static int decide(ref_model *m, double lMHr, uint64_t seed, int c, uint64_t iter, int sid, int t, int k) {
  double u, u2;
  if (isnan(lMHr)) return 0;
  if (lMHr > 0) return 1;
  ref_unif2(seed, (uint32_t)(m->first_chain + c), iter, sid, t, k, &u, &u2);
  return u < exp(lMHr);
}

/* ------------------------------------------------------------------ samplers */
/* column Cholesky, lower L with A = L L^T, returned as upper U = L^T (column-major), fixed loop order */
static void chol_upper(const double *A, double *U, int d) {
  int i, j, k;
  for (i = 0; i < d * d; i++) U[i] = 0.0;
  for (j = 0; j < d; j++) {
    double s = A[j + j * d];
    for (k = 0; k < j; k++) s -= U[k + j * d] * U[k + j * d];
    s = sqrt(s);
    U[j + j * d] = s;
    for (i = j + 1; i < d; i++) {
      double v = A[j + i * d];
      for (k = 0; k < j; k++) v -= U[k + j * d] * U[k + i * d];
      U[j + i * d] = v / s;
    }
  }
}

static void sampler_reset(const nimb200_sampler *sp, SState *s, int adaptInterval) {
  s->scale = s->scale0; s->timesRan = s->timesAccepted = s->timesAdapted = 0; s->gamma1 = 0; s->sumJumps = 0;
  if (sp->type == NIMB_SAMPLER_RW_BLOCK) {
    int d = sp->d, i;
    memcpy(s->propCov, s->propCov0, sizeof(double) * d * d);
    chol_upper(s->propCov, s->chol, d);
    for (i = 0; i < d * d; i++) s->cholScale[i] = s->chol[i] * s->scale;
    memset(s->empirSamp, 0, sizeof(double) * adaptInterval * d);
  }
}

/* sampler_RW::run, MCMC_samplers.R:352-386 + adaptiveProcedure :388-420 */
static void run_RW(ref_model *m, int c, int sid, int t, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  SState *s = m->ss[sid] + (size_t)t * m->C + c;
  double *st = m->state + (size_t)c * m->t.n_slots;
  int slot = sp->target_slot + t, jump;
  double cur = st[slot], propLogScale = 0, prop, z, z1, logMHR;
  ref_normal2(seed, (uint32_t)(m->first_chain + c), iter, sid, t, 0, &z, &z1);
  if (sp->log_scale) {
    propLogScale = 0.0 + s->scale * z;
    prop = cur * exp(propLogScale);
  } else
    prop = cur + s->scale * z;
  if (sp->reflective) {
    while (prop < sp->lower || prop > sp->upper) {
      if (prop < sp->lower) prop = 2 * sp->lower - prop;
      if (prop > sp->upper) prop = 2 * sp->upper - prop;
    }
  }
  st[slot] = prop;
  logMHR = calculateDiff_deps(m, c, sp, t, 0);
  if (logMHR == -INFINITY) {
    jump = 0;
    copy_targets(m, c, sp, t, 0);
    copy_lp(m, c, sp, t, 0, 0);
  } else {
    logMHR = logMHR + calculateDiff_deps(m, c, sp, t, 1) + propLogScale;
    jump = decide(m, logMHR, seed, c, iter, sid, t, 1);
    copy_targets(m, c, sp, t, jump);
    copy_lp(m, c, sp, t, 0, jump);
    copy_lp(m, c, sp, t, 1, jump);
  }
  if (sp->adaptive) {
    s->timesRan++;
    if (jump) s->timesAccepted++;
    if (s->timesRan % sp->adapt_interval == 0) {
      double acceptanceRate = (double)s->timesAccepted / (double)s->timesRan, gamma2, adaptFactor;
      s->timesAdapted++;
      s->gamma1 = 1 / pow((double)s->timesAdapted + 3, sp->adapt_factor_exponent);
      gamma2 = 10 * s->gamma1;
      adaptFactor = exp(gamma2 * (acceptanceRate - 0.44));
      s->scale = s->scale * adaptFactor;
      if (sp->reflective && s->scale >= 0.5 * (sp->upper - sp->lower)) s->scale = 0.5 * (sp->upper - sp->lower);
      s->timesRan = 0;
      s->timesAccepted = 0;
    }
  }
}

/* sampler_RW_block::run, MCMC_samplers.R:688-714; proposal rmnorm_chol dists.cpp:1132-1164 (dtrmv 'U','T','N') */
static void run_RW_block(ref_model *m, int c, int sid, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  SState *s = m->ss[sid] + c;
  double *st = m->state + (size_t)c * m->t.n_slots;
  int d = sp->d, s0 = sp->target_slot, tr, i, j, jump;
  int npair = (d + 1) / 2;
  double x[NIMB200_MAX_BLOCK_D];
  static const double optAR[5] = {0.44, 0.35, 0.32, 0.25, 0.234};
  for (tr = 0; tr < sp->tries; tr++) {
    double lpD;
    for (j = 0; j < npair; j++) {
      double z0, z1;
      ref_normal2(seed, (uint32_t)(m->first_chain + c), iter, sid, 0, tr * (npair + 1) + j, &z0, &z1);
      x[2 * j] = z0;
      if (2 * j + 1 < d) x[2 * j + 1] = z1;
    }
    for (j = d - 1; j >= 0; j--) { /* x := U^T x, reference-BLAS loop order */
      double temp = x[j] * s->cholScale[j + j * d];
      for (i = j - 1; i >= 0; i--) temp += s->cholScale[i + j * d] * x[i];
      x[j] = temp;
    }
    for (j = 0; j < d; j++) x[j] += st[s0 + j];
    for (j = 0; j < d; j++) st[s0 + j] = x[j];
    lpD = calculateDiff_deps(m, c, sp, 0, 0);
    if (lpD == -INFINITY) {
      jump = 0;
      copy_targets(m, c, sp, 0, 0);
      copy_lp(m, c, sp, 0, 0, 0);
    } else {
      lpD = lpD + calculateDiff_deps(m, c, sp, 0, 1);
      jump = decide(m, lpD, seed, c, iter, sid, 0, tr * (npair + 1) + npair);
      copy_targets(m, c, sp, 0, jump);
      copy_lp(m, c, sp, 0, 0, jump);
      copy_lp(m, c, sp, 0, 1, jump);
    }
    if (sp->adaptive) { /* adaptiveProcedure, MCMC_samplers.R:716-760 */
      int AI = sp->adapt_interval;
      s->timesRan++;
      if (jump) s->timesAccepted++;
      if (!sp->adapt_scale_only)
        for (j = 0; j < d; j++) s->empirSamp[(s->timesRan - 1) + (size_t)j * AI] = st[s0 + j];
      if (s->timesRan % AI == 0) {
        double acceptanceRate = (double)s->timesAccepted / (double)s->timesRan, gamma2, adaptFactor;
        s->timesAdapted++;
        s->gamma1 = 1 / pow((double)s->timesAdapted + 3, sp->adapt_factor_exponent);
        gamma2 = 10 * s->gamma1;
        adaptFactor = exp(gamma2 * (acceptanceRate - optAR[d > 5 ? 4 : d - 1]));
        s->scale = s->scale * adaptFactor;
        if (!sp->adapt_scale_only) {
          int a, b, r;
          for (j = 0; j < d; j++) {
            double mu = 0;
            for (r = 0; r < AI; r++) mu += s->empirSamp[r + (size_t)j * AI];
            mu = mu / AI;
            for (r = 0; r < AI; r++) s->empirSamp[r + (size_t)j * AI] -= mu;
          }
          for (a = 0; a < d; a++)
            for (b = 0; b < d; b++) {
              double acc = 0, ec;
              for (r = 0; r < AI; r++) acc += s->empirSamp[r + (size_t)a * AI] * s->empirSamp[r + (size_t)b * AI];
              ec = acc / (double)(s->timesRan - 1);
              s->propCov[a + b * d] = s->propCov[a + b * d] + s->gamma1 * (ec - s->propCov[a + b * d]);
            }
          chol_upper(s->propCov, s->chol, d);
        }
        for (i = 0; i < d * d; i++) s->cholScale[i] = s->chol[i] * s->scale;
        s->timesRan = 0;
        s->timesAccepted = 0;
      }
    }
  }
}

/* sampler_binary::run, MCMC_samplers.R:127-151 */
static void run_binary(ref_model *m, int c, int sid, int t, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  double *st = m->state + (size_t)c * m->t.n_slots;
  int slot = sp->target_slot + t, jump;
  double currentLogProb = getLogProb_deps(m, c, sp, t, -1), otherLogProbPrior, otherLogProb, acceptanceProb, u, u2;
  st[slot] = 1 - st[slot];
  otherLogProbPrior = calculate_deps(m, c, sp, t, 0);
  if (otherLogProbPrior == -INFINITY) otherLogProb = otherLogProbPrior;
  else otherLogProb = otherLogProbPrior + calculate_deps(m, c, sp, t, 1);
  acceptanceProb = 1 / (exp(currentLogProb - otherLogProb) + 1);
  ref_unif2(seed, (uint32_t)(m->first_chain + c), iter, sid, t, 0, &u, &u2);
  jump = (!isnan(acceptanceProb)) & (u < acceptanceProb);
  copy_targets(m, c, sp, t, jump);
  copy_lp(m, c, sp, t, 0, jump);
  copy_lp(m, c, sp, t, 1, jump);
}

/* generated conjugate sampler run(): MCMC_conjugacy.R:651-692 with :721-1034 */
static void run_conjugate(ref_model *m, int c, int sid, int t, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  double *st = m->state + (size_t)c * m->t.n_slots;
  int slot = sp->target_slot + t, q, pass;
  const nimb200_dep *tp = m->t.deps + sp->dep_off; /* stage-0 entry: the target's own declaration */
  int tnode = sel_of(m, tp, t).lo;
  int gam = sp->type == NIMB_SAMPLER_CONJ_GAMMA, bet = sp->type == NIMB_SAMPLER_CONJ_BETA;
  double prior0 = get_param(m, st, tp->decl, tnode, 0);                                    /* prior_shape | prior_mean | prior_shape1 */
  double prior1 = get_param(m, st, tp->decl, tnode, bet ? 1 : (gam ? NIMB_PARAM_RATE : NIMB_PARAM_TAU)); /* prior_rate | prior_tau | prior_shape2 */
  double contrib0 = 0, contrib1 = 0, saved_target = st[slot], newValue;
  /* count dependents to size the dep_* work arrays the generated code keeps */
  int ndeps = 0;
  FOR_DEP_NODES(m, sp, t, 1, { (void)node; (void)di; ndeps++; });
  double *work = (double *)malloc(sizeof(double) * 4 * (ndeps + 1));
  double *dval = work, *dpar = work + ndeps, *doff = work + 2 * ndeps, *dcoef = work + 3 * ndeps;
  q = 0;
  FOR_DEP_NODES(m, sp, t, 1, {
    const nimb200_decl *dd = m->t.decl + di;
    dval[q] = node_value(m, st, dd, node);
    /* neededParamsForPosterior: dnorm dependents need tau (dnorm prior) or mean (dgamma prior on tau);
       dgamma dependents need shape; dpois none */
    if (dd->dist == NIMB_DNORM || dd->dist == NIMB_DLNORM) dpar[q] = gam ? get_param(m, st, di, node, 0) : get_param(m, st, di, node, NIMB_PARAM_TAU);
    else if (dd->dist == NIMB_DGAMMA) dpar[q] = get_param(m, st, di, node, 0);
    else if (dd->dist == NIMB_DBINOM) dpar[q] = get_param(m, st, di, node, 0); /* size (1 for dbern) */
    else dpar[q] = 0;
    doff[q] = 0; dcoef[q] = 1;
    q++;
  });
  for (pass = 0; pass < 2; pass++) { /* target <- 0 (offsets), then target <- 1 (coefficients): :766-820 */
    int need = 0;
    for (q = 0; q < sp->n_dep; q++) {
      const nimb200_dep *dp = m->t.deps + sp->dep_off + q;
      if (dp->stage != 1) continue;
      if (pass == 0 && (dp->link == NIMB_LINK_ADDITIVE || dp->link == NIMB_LINK_LINEAR)) need = 1;
      if (pass == 1 && (dp->link == NIMB_LINK_MULTIPLICATIVE || dp->link == NIMB_LINK_LINEAR)) need = 1;
    }
    if (!need) continue;
    st[slot] = pass == 0 ? 0.0 : 1.0;
    q = 0;
    FOR_DEP_NODES(m, sp, t, 1, {
      if (pass == 0 && (dp->link == NIMB_LINK_ADDITIVE || dp->link == NIMB_LINK_LINEAR))
        doff[q] = get_param(m, st, di, node, dp->param);
      if (pass == 1 && dp->link == NIMB_LINK_LINEAR) dcoef[q] = get_param(m, st, di, node, dp->param) - doff[q];
      if (pass == 1 && dp->link == NIMB_LINK_MULTIPLICATIVE) dcoef[q] = get_param(m, st, di, node, dp->param);
      q++;
    });
  }
  q = 0;
  FOR_DEP_NODES(m, sp, t, 1, {
    const nimb200_decl *dd = m->t.decl + di;
    double value = dval[q], coeff = dcoef[q], offset = doff[q];
    if (bet) { /* MCMC_conjugacy.R:4-8: dbern value, 1 - value; dbin value, size - value */
      contrib0 += value; contrib1 += dpar[q] - value;
    } else if (gam) { /* MCMC_conjugacy.R:80-93 */
      if (dd->dist == NIMB_DPOIS) { contrib0 += value; contrib1 += coeff; }
      else if (dd->dist == NIMB_DGAMMA) { contrib0 += dpar[q]; contrib1 += coeff * value; }
      else if (dd->dist == NIMB_DEXP) { contrib0 += 1; contrib1 += coeff * value; }
      else if (dd->dist == NIMB_DNORM) { contrib0 += 1. / 2; contrib1 += coeff / 2 * ((value - dpar[q]) * (value - dpar[q])); }
      else if (dd->dist == NIMB_DLNORM) { contrib0 += 1. / 2; contrib1 += coeff / 2 * ((log(value) - dpar[q]) * (log(value) - dpar[q])); }
    } else { /* :109-115; a dlnorm dependent contributes through log(value) */
      double tau = dpar[q];
      if (dd->dist == NIMB_DLNORM) value = log(value);
      contrib0 += coeff * (value - offset) * tau;
      contrib1 += coeff * coeff * tau;
    }
    q++;
  });
  free(work);
  (void)saved_target;
  if (bet) {
    newValue = ref_rbeta(seed, (uint32_t)(m->first_chain + c), iter, sid, t, prior0 + contrib0, prior1 + contrib1);
  } else if (gam) {
    double shape = prior0 + contrib0, scale = 1 / (prior1 + contrib1);
    newValue = ref_rgamma(seed, (uint32_t)(m->first_chain + c), iter, sid, t, shape) * scale;
  } else {
    double mean = (prior0 * prior1 + contrib0) / (prior1 + contrib1), sd = pow(prior1 + contrib1, -0.5), z, z1;
    ref_normal2(seed, (uint32_t)(m->first_chain + c), iter, sid, t, 0, &z, &z1);
    newValue = mean + sd * z;
  }
  st[slot] = newValue;
  calculate_deps(m, c, sp, t, -1);
  copy_targets(m, c, sp, t, 1);
  copy_lp(m, c, sp, t, -1, 1);
}

/* sampler_slice::run (MCMC_samplers.R:893-938) with setAndCalculateTarget (:941-949) and adaptiveProcedure (:950-965).
 * Uniforms from the sampler's counter: call 0 = (auxiliary exponential, interval offset), call 1 = (split of maxSteps, first
 * candidate), call 2 + j = j-th shrinkage candidate.  getLogProb(calcNodes) is taken as (target) + (dependents). */
static double slice_set_and_calc(ref_model *m, int c, const nimb200_sampler *sp, int t, int discrete, double value) {
  double *st = m->state + (size_t)c * m->t.n_slots, lp;
  st[sp->target_slot + t] = discrete ? floor(value) : value;
  lp = calculate_deps(m, c, sp, t, 0);
  if (lp == -INFINITY) return -INFINITY;
  return lp + calculate_deps(m, c, sp, t, 1);
}
static void run_slice(ref_model *m, int c, int sid, int t, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  double *st = m->state + (size_t)c * m->t.n_slots;
  SState *S = m->ss[sid] + (size_t)t * m->C + c;
  int slot = sp->target_slot + t;
  int tdist = m->t.decl[m->t.deps[sp->dep_off].decl].dist;
  int discrete = tdist == NIMB_DPOIS || tdist == NIMB_DBINOM || tdist == NIMB_DCAT;
  uint32_t chain = (uint32_t)(m->first_chain + c);
  const double eps = 1e-15, x0 = st[slot], width = S->scale;
  double ua, ub, uc, ud, ue, uf, u, L, R, lp, x1, cur0;
  int maxStepsL, maxStepsR, nc = 0;
  ref_unif2(seed, chain, iter, sid, t, 0, &ua, &ub);
  ref_unif2(seed, chain, iter, sid, t, 1, &uc, &ud);
  cur0 = getLogProb_deps(m, c, sp, t, 0);
  u = (cur0 == -INFINITY ? -INFINITY : cur0 + getLogProb_deps(m, c, sp, t, 1)) - (-ref_det_log(ua));
  L = x0 - ub * width; R = L + width;
  maxStepsL = (int)floor(uc * (double)sp->tries); maxStepsR = sp->tries - 1 - maxStepsL;
  lp = slice_set_and_calc(m, c, sp, t, discrete, L);
  while (maxStepsL > 0 && !isnan(lp) && lp >= u) { L = L - width; lp = slice_set_and_calc(m, c, sp, t, discrete, L); maxStepsL--; }
  lp = slice_set_and_calc(m, c, sp, t, discrete, R);
  while (maxStepsR > 0 && !isnan(lp) && lp >= u) { R = R + width; lp = slice_set_and_calc(m, c, sp, t, discrete, R); maxStepsR--; }
  x1 = L + ud * (R - L);
  lp = slice_set_and_calc(m, c, sp, t, discrete, x1);
  while ((isnan(lp) || lp < u) && (R - L) / (fabs(R) + fabs(L) + eps) > eps && nc < sp->pad) {
    if (x1 < x0) L = x1; else R = x1;
    ref_unif2(seed, chain, iter, sid, t, 2 + nc, &ue, &uf);
    x1 = L + ue * (R - L);
    lp = slice_set_and_calc(m, c, sp, t, discrete, x1);
    nc++;
  }
  if ((R - L) / (fabs(R) + fabs(L) + eps) <= eps || nc == sp->pad) {
    copy_targets(m, c, sp, t, 0);
    copy_lp(m, c, sp, t, -1, 0);
    return;
  }
  copy_targets(m, c, sp, t, 1);
  copy_lp(m, c, sp, t, -1, 1);
  if (sp->adaptive) {
    S->timesRan++;
    S->sumJumps += fabs(x1 - x0);
    if (S->timesRan % sp->adapt_interval == 0) {
      double af = 1.0, meanJump;
      long i;
      for (i = 0; i < S->timesAdapted; i++) af = af * 0.75;
      meanJump = S->sumJumps / (double)S->timesRan;
      S->scale = width + (2 * meanJump - width) * af;
      S->timesAdapted++;
      S->timesRan = 0; S->sumJumps = 0;
    }
  }
}

/* sampler_categorical::run (MCMC_samplers.R:187-216); rcat: dists.cpp:936-961 with the uniform from the sampler's counter */
static int run_categorical(ref_model *m, int c, int sid, int t, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  double *st = m->state + (size_t)c * m->t.n_slots;
  int slot = sp->target_slot + t, K = m->t.decl[m->t.deps[sp->dep_off].decl].fam_arg[0], i, value = 1;
  double lps[NIMB200_MAX_CATEGORIES], maxLP = -INFINITY, sum = 0.0, u, u2, cum;
  if (K > NIMB200_MAX_CATEGORIES) return fail("categorical: too many categories");
  for (i = 1; i <= K; i++) {
    double lp = slice_set_and_calc(m, c, sp, t, 0, (double)i);
    if (isnan(lp)) lp = -INFINITY;
    lps[i - 1] = lp;
    maxLP = fmax(maxLP, lp);
  }
  for (i = 0; i < K; i++) { lps[i] = ref_det_exp(lps[i] - maxLP); sum += lps[i]; }
  if (isnan(sum)) { copy_targets(m, c, sp, t, 0); calculate_deps(m, c, sp, t, -1); return 0; }
  ref_unif2(seed, (uint32_t)(m->first_chain + c), iter, sid, t, 0, &u, &u2);
  u = u * sum;
  cum = lps[0];
  while (u > cum && value < K) { cum += lps[value]; value++; }
  st[slot] = (double)value;
  calculate_deps(m, c, sp, t, -1);
  copy_targets(m, c, sp, t, 1);
  copy_lp(m, c, sp, t, -1, 1);
  return 0;
}

/* sampler_posterior_predictive::run (MCMC_samplers.R:97-101) for a terminal node: model$simulate(target), calculate, copy.
 * The variate generators are the counter-based ones of draws_ref.c (call 0.. of the sampler's counter), not R's. */
static int run_posterior_predictive(ref_model *m, int c, int sid, int t, uint64_t seed, uint64_t iter) {
  const nimb200_sampler *sp = m->t.samplers + sid;
  double *st = m->state + (size_t)c * m->t.n_slots;
  int slot = sp->target_slot + t;
  const nimb200_dep *tp = m->t.deps + sp->dep_off;
  int tnode = sel_of(m, tp, t).lo;
  const nimb200_decl *d = m->t.decl + tp->decl;
  uint32_t chain = (uint32_t)(m->first_chain + c);
  double p0 = get_param(m, st, tp->decl, tnode, 0), p1 = d->nparam > 1 ? get_param(m, st, tp->decl, tnode, 1) : 0.0;
  double u, u2, nv;
  switch (d->dist) {
    case NIMB_DNORM: ref_normal2(seed, chain, iter, sid, t, 0, &u, &u2); nv = p0 + p1 * u; break;
    case NIMB_DGAMMA: nv = ref_rgamma(seed, chain, iter, sid, t, p0) * p1; break;
    case NIMB_DBETA: nv = ref_rbeta(seed, chain, iter, sid, t, p0, p1); break;
    case NIMB_DEXP: ref_unif2(seed, chain, iter, sid, t, 0, &u, &u2); nv = -ref_det_log(u) / p0; break;
    case NIMB_DLNORM: ref_normal2(seed, chain, iter, sid, t, 0, &u, &u2); nv = ref_det_exp(p0 + p1 * u); break;
    case NIMB_DLOGIS: ref_unif2(seed, chain, iter, sid, t, 0, &u, &u2); nv = p0 + p1 * (ref_det_log(u) - ref_det_log(1.0 - u)); break;
    case NIMB_DBINOM: nv = ref_rbinom(seed, chain, iter, sid, t, p0, p1); break;
    case NIMB_DPOIS: nv = ref_rpois(seed, chain, iter, sid, t, p0); break;
    default: return fail("posterior_predictive: unsupported distribution");
  }
  st[slot] = nv;
  calculate_deps(m, c, sp, t, -1);
  copy_targets(m, c, sp, t, 1);
  copy_lp(m, c, sp, t, -1, 1);
  return 0;
}

/* ------------------------------------------------------------------ model life cycle */
int ref_model_create(const nimb200_table *t, int32_t nchains, int32_t first_chain, ref_model **out) {
  ref_model *m;
  int i, s;
  if (t->version != NIMB200_TABLE_VERSION) return fail("table version mismatch");
  m = (ref_model *)calloc(1, sizeof *m);
  m->t = *t;
  m->C = nchains; m->first_chain = first_chain;
  m->darr = (double **)calloc(t->n_darr + 1, sizeof(double *));
  for (i = 0; i < t->n_darr; i++) {
    m->darr[i] = (double *)malloc(sizeof(double) * (t->darr_len[i] + 1));
    memcpy(m->darr[i], t->darr[i], sizeof(double) * t->darr_len[i]);
  }
  m->iarr = (int32_t **)calloc(t->n_iarr + 1, sizeof(int32_t *));
  for (i = 0; i < t->n_iarr; i++) {
    m->iarr[i] = (int32_t *)malloc(sizeof(int32_t) * (t->iarr_len[i] + 1));
    memcpy(m->iarr[i], t->iarr[i], sizeof(int32_t) * t->iarr_len[i]);
  }
#define DUP(field, type, n) { type *p_ = (type *)malloc(sizeof(type) * ((n) + 1)); memcpy(p_, t->field, sizeof(type) * (n)); m->t.field = p_; }
  DUP(ops, nimb200_op, t->n_ops) DUP(decl, nimb200_decl, t->n_decl) DUP(samplers, nimb200_sampler, t->n_samplers)
  DUP(deps, nimb200_dep, t->n_deps) DUP(monitor_slots, int32_t, t->n_monitors)
  m->state = (double *)calloc((size_t)nchains * t->n_slots + 1, sizeof(double));
  m->saved = (double *)calloc((size_t)nchains * t->n_slots + 1, sizeof(double));
  m->lp = (double **)calloc(t->n_decl + 1, sizeof(double *));
  m->lp_saved = (double **)calloc(t->n_decl + 1, sizeof(double *));
  for (i = 0; i < t->n_decl; i++) {
    m->lp[i] = (double *)calloc((size_t)nchains * t->decl[i].n + 1, sizeof(double));
    m->lp_saved[i] = (double *)calloc((size_t)nchains * t->decl[i].n + 1, sizeof(double));
  }
  m->ss = (SState **)calloc(t->n_samplers + 1, sizeof(SState *));
  for (s = 0; s < t->n_samplers; s++) {
    const nimb200_sampler *sp = t->samplers + s;
    int T = sp->type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp->n_targets, q;
    m->ss[s] = (SState *)calloc((size_t)T * nchains, sizeof(SState));
    for (q = 0; q < T * nchains; q++) {
      SState *st = m->ss[s] + q;
      st->scale = st->scale0 = sp->scale;
      if (sp->type == NIMB_SAMPLER_RW_BLOCK) {
        int d = sp->d, a;
        if (d > NIMB200_MAX_BLOCK_D) return fail("RW_block dimension too large");
        st->propCov = (double *)calloc(d * d, sizeof(double));
        st->propCov0 = (double *)calloc(d * d, sizeof(double));
        st->chol = (double *)calloc(d * d, sizeof(double));
        st->cholScale = (double *)calloc(d * d, sizeof(double));
        st->empirSamp = (double *)calloc((size_t)sp->adapt_interval * d, sizeof(double));
        if (sp->propcov_darr >= 0) memcpy(st->propCov0, t->darr[sp->propcov_darr], sizeof(double) * d * d);
        else for (a = 0; a < d; a++) st->propCov0[a + a * d] = 1.0;
      }
      sampler_reset(sp, st, sp->adapt_interval);
    }
  }
  *out = m;
  return 0;
}

int ref_model_free(ref_model *m) {
  int i, s;
  if (!m) return 0;
  for (i = 0; i < m->t.n_darr; i++) free(m->darr[i]);
  for (i = 0; i < m->t.n_iarr; i++) free(m->iarr[i]);
  for (i = 0; i < m->t.n_decl; i++) { free(m->lp[i]); free(m->lp_saved[i]); }
  for (s = 0; s < m->t.n_samplers; s++) {
    const nimb200_sampler *sp = m->t.samplers + s;
    int T = sp->type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp->n_targets, q;
    if (sp->type == NIMB_SAMPLER_RW_BLOCK)
      for (q = 0; q < T * m->C; q++) {
        SState *st = m->ss[s] + q;
        free(st->propCov); free(st->propCov0); free(st->chol); free(st->cholScale); free(st->empirSamp);
      }
    free(m->ss[s]);
  }
  free((void *)m->t.ops); free((void *)m->t.decl); free((void *)m->t.samplers); free((void *)m->t.deps);
  free((void *)m->t.monitor_slots);
  free(m->darr); free(m->iarr); free(m->lp); free(m->lp_saved); free(m->ss);
  free(m->state); free(m->saved); free(m->samples);
  free(m);
  return 0;
}

int ref_set_values(ref_model *m, const double *v) { memcpy(m->state, v, sizeof(double) * (size_t)m->C * m->t.n_slots); return 0; }
int ref_get_values(ref_model *m, double *v) { memcpy(v, m->state, sizeof(double) * (size_t)m->C * m->t.n_slots); return 0; }

int ref_calculate_nodes(ref_model *m, int32_t di, int32_t lo, int32_t hi, double *out) {
  int c, i;
  if (di < 0 || di >= m->t.n_decl || lo < 0 || hi > m->t.decl[di].n || lo > hi) return fail("bad node range");
  for (c = 0; c < m->C; c++) {
    const double *st = m->state + (size_t)c * m->t.n_slots;
    double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n;
    for (i = lo; i < hi; i++) { lp[i] = calc_node(m, st, di, i); out[(size_t)(i - lo) * m->C + c] = lp[i]; }
  }
  return 0;
}
int ref_calculate(ref_model *m, int32_t di, int32_t lo, int32_t hi, double *out) {
  int c, i;
  if (di < 0 || di >= m->t.n_decl || lo < 0 || hi > m->t.decl[di].n || lo > hi) return fail("bad node range");
  for (c = 0; c < m->C; c++) {
    const double *st = m->state + (size_t)c * m->t.n_slots;
    double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n, ans = 0;
    for (i = lo; i < hi; i++) { lp[i] = calc_node(m, st, di, i); ans += lp[i]; }
    out[c] = ans;
  }
  return 0;
}
int ref_calculateDiff(ref_model *m, int32_t di, double *out) {
  int c, i;
  if (di < 0 || di >= m->t.n_decl) return fail("bad declaration");
  for (c = 0; c < m->C; c++) {
    const double *st = m->state + (size_t)c * m->t.n_slots;
    double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n, ans = 0;
    for (i = 0; i < m->t.decl[di].n; i++) { double nw = calc_node(m, st, di, i); ans += nw - lp[i]; lp[i] = nw; }
    out[c] = ans;
  }
  return 0;
}
int ref_getLogProb(ref_model *m, int32_t di, double *out) {
  int c, i;
  if (di < 0 || di >= m->t.n_decl) return fail("bad declaration");
  for (c = 0; c < m->C; c++) {
    const double *lp = m->lp[di] + (size_t)c * m->t.decl[di].n;
    double ans = 0;
    for (i = 0; i < m->t.decl[di].n; i++) ans += lp[i];
    out[c] = ans;
  }
  return 0;
}

/* MCMC::run, MCMC_build.R:217-321; chains one after another, MCMC_run.R:127-153.
 * Split in three so that bench.py can give disjoint chain ranges to host threads (the manual's
 * one-process-per-core pattern, UserManual/src/chapter_MCMC.Rmd:1223-1229): prepare (argument checks,
 * mvSamples resize), run_range (the chain loop itself), finish (row / iteration bookkeeping). */
int ref_mcmc_prepare(ref_model *m, int64_t niter, int64_t nburnin, int32_t thin, int32_t reset) {
  int c;
  int64_t addrows;
  const nimb200_table *T = &m->t;
  if (niter < 0) return fail("cannot specify niter < 0");
  if (nburnin < 0) return fail("cannot specify nburnin < 0");
  if (nburnin > niter) return fail("cannot specify nburnin > niter");
  if (thin < 1) return fail("cannot use thin < 1");
  if (!reset && nburnin != 0) return fail("cannot specify nburnin when using reset = FALSE.");
  if (reset) {
    m->nrows = 0; m->iter_done = 0;
    if (m->waic_acc) {
      int di;
      m->waic_count = 0;
      for (di = 0; di < T->n_decl; di++)
        if (m->waic_acc[di]) memset(m->waic_acc[di], 0, sizeof(double) * 4 * T->decl[di].n);
    }
  }
  addrows = (niter - nburnin) / thin;
  if (m->nrows + addrows > m->cap || !m->samples) {
    int64_t ncap = m->nrows + addrows;
    double *ns = (double *)calloc((size_t)m->C * (ncap + 1) * (T->n_monitors + 1), sizeof(double));
    for (c = 0; c < m->C && m->samples; c++)
      memcpy(ns + (size_t)c * ncap * T->n_monitors, m->samples + (size_t)c * m->cap * T->n_monitors,
             sizeof(double) * m->nrows * T->n_monitors);
    free(m->samples);
    m->samples = ns; m->cap = ncap;
  }
  return 0;
}

int ref_mcmc_run_range(ref_model *m, int64_t niter, int64_t nburnin, int32_t thin, int32_t reset, uint64_t seed,
                       int32_t chain_lo, int32_t chain_hi) {
  int c, s, t, di, i;
  int64_t iter;
  const nimb200_table *T = &m->t;
  for (c = chain_lo; c < chain_hi; c++) {
    double *st = m->state + (size_t)c * T->n_slots;
    int64_t row = m->nrows;
    /* initializeModel: every stochastic node calculated once */
    for (di = 0; di < T->n_decl; di++) {
      double *lp = m->lp[di] + (size_t)c * T->decl[di].n;
      for (i = 0; i < T->decl[di].n; i++) lp[i] = calc_node(m, st, di, i);
      memcpy(m->lp_saved[di] + (size_t)c * T->decl[di].n, lp, sizeof(double) * T->decl[di].n);
    }
    memcpy(m->saved + (size_t)c * T->n_slots, st, sizeof(double) * T->n_slots);
    if (reset)
      for (s = 0; s < T->n_samplers; s++) {
        const nimb200_sampler *sp = T->samplers + s;
        int TT = sp->type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp->n_targets;
        for (t = 0; t < TT; t++) sampler_reset(sp, m->ss[s] + (size_t)t * m->C + c, sp->adapt_interval);
      }
    for (iter = 1; iter <= niter; iter++) {
      uint64_t git = (uint64_t)(m->iter_done + iter);
      for (s = 0; s < T->n_samplers; s++) {
        const nimb200_sampler *sp = T->samplers + s;
        switch (sp->type) {
          case NIMB_SAMPLER_RW: for (t = 0; t < sp->n_targets; t++) run_RW(m, c, s, t, seed, git); break;
          case NIMB_SAMPLER_RW_BLOCK: run_RW_block(m, c, s, seed, git); break;
          case NIMB_SAMPLER_BINARY: for (t = 0; t < sp->n_targets; t++) run_binary(m, c, s, t, seed, git); break;
          case NIMB_SAMPLER_CONJ_GAMMA:
          case NIMB_SAMPLER_CONJ_NORMAL:
          case NIMB_SAMPLER_CONJ_BETA: for (t = 0; t < sp->n_targets; t++) run_conjugate(m, c, s, t, seed, git); break;
          case NIMB_SAMPLER_SLICE: for (t = 0; t < sp->n_targets; t++) run_slice(m, c, s, t, seed, git); break;
          case NIMB_SAMPLER_CATEGORICAL:
            for (t = 0; t < sp->n_targets; t++) if (run_categorical(m, c, s, t, seed, git)) return 1;
            break;
          case NIMB_SAMPLER_POSTERIOR_PREDICTIVE:
            for (t = 0; t < sp->n_targets; t++) if (run_posterior_predictive(m, c, s, t, seed, git)) return 1;
            break;
          default: return fail("unknown sampler type");
        }
      }
      if (iter > nburnin && m->waic) { /* waicFun$updateStats(), MCMC_build.R:304-310, MCMC_WAIC.R:209-268 (conditional, no groups) */
        double cnt = (double)(++m->waic_count);
        for (di = 0; di < T->n_decl; di++) {
          int n = T->decl[di].n, node;
          double *acc = m->waic_acc[di];
          const double *lpv;
          if (!acc) continue;
          lpv = m->lp[di] + (size_t)c * n;
          for (node = 0; node < n; node++) {
            double lp = lpv[node], d1, d2;
            if (m->waic_count == 1) { acc[node] = lp; acc[n + node] = 1.0; }
            else if (lp > acc[node]) { acc[n + node] = acc[n + node] * exp(-lp + acc[node]) + 1.0; acc[node] = lp; }
            else acc[n + node] = acc[n + node] + exp(lp - acc[node]);
            d1 = lp - acc[2 * n + node];
            acc[2 * n + node] = acc[2 * n + node] + d1 / cnt;
            d2 = lp - acc[2 * n + node];
            acc[3 * n + node] = acc[3 * n + node] + d1 * d2;
          }
        }
      }
      if (iter > nburnin) {
        int64_t sampleNumber = iter - nburnin;
        if (sampleNumber % thin == 0) {
          double *dst = m->samples + ((size_t)c * m->cap + row) * T->n_monitors;
          for (i = 0; i < T->n_monitors; i++) dst[i] = st[T->monitor_slots[i]];
          row++;
        }
      }
    }
  }
  __atomic_fetch_add(&m->evals, tl_evals, __ATOMIC_RELAXED);
  tl_evals = 0;
  return 0;
}
int ref_mcmc_finish(ref_model *m, int64_t niter, int64_t nburnin, int32_t thin) {
  m->nrows += (niter - nburnin) / thin;
  m->iter_done += niter;
  return 0;
}
int ref_mcmc_run(ref_model *m, int64_t niter, int64_t nburnin, int32_t thin, int32_t reset, uint64_t seed) {
  int rc = ref_mcmc_prepare(m, niter, nburnin, thin, reset);
  if (rc) return rc;
  rc = ref_mcmc_run_range(m, niter, nburnin, thin, reset, seed, 0, m->C);
  if (rc) return rc;
  return ref_mcmc_finish(m, niter, nburnin, thin);
}

int ref_mcmc_samples(ref_model *m, double *out, int64_t *nrows_out) {
  int c;
  if (nrows_out) *nrows_out = m->nrows;
  if (!out) return 0;
  for (c = 0; c < m->C; c++)
    memcpy(out + (size_t)c * m->nrows * m->t.n_monitors, m->samples + (size_t)c * m->cap * m->t.n_monitors,
           sizeof(double) * m->nrows * m->t.n_monitors);
  return 0;
}

int ref_sampler_state(ref_model *m, int32_t s, int32_t what, double *out, int64_t cap) {
  const nimb200_sampler *sp;
  int T, t, c;
  if (s < 0 || s >= m->t.n_samplers) return fail("bad sampler index");
  sp = m->t.samplers + s;
  T = sp->type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp->n_targets;
  if (what == 0 || what == 1) {
    if (cap < (int64_t)T * m->C) return fail("output too small");
    for (t = 0; t < T; t++)
      for (c = 0; c < m->C; c++) {
        SState *st = m->ss[s] + (size_t)t * m->C + c;
        out[(size_t)t * m->C + c] = what == 0 ? st->scale : (double)st->timesAdapted;
      }
    return 0;
  }
  if (what == 2 && sp->type == NIMB_SAMPLER_RW_BLOCK) {
    int d = sp->d, q;
    if (cap < (int64_t)d * d * m->C) return fail("output too small");
    for (q = 0; q < d * d; q++)
      for (c = 0; c < m->C; c++) out[(size_t)q * m->C + c] = m->ss[s][c].propCov[q];
    return 0;
  }
  return fail("bad state selector");
}

uint64_t ref_eval_count(ref_model *m) { uint64_t v = m->evals + tl_evals; return v; }

int ref_waic_enable(ref_model *m, int32_t enable) {
  int di;
  if (enable && !m->waic_acc) {
    m->waic_acc = (double **)calloc(m->t.n_decl + 1, sizeof(double *));
    for (di = 0; di < m->t.n_decl; di++)
      if (m->t.decl[di].value_kind == 0 && m->t.decl[di].n > 0) m->waic_acc[di] = (double *)calloc((size_t)4 * m->t.decl[di].n, sizeof(double));
    m->waic_count = 0;
  }
  m->waic = enable != 0;
  return 0;
}
int ref_waic_get(ref_model *m, double *out4) {
  int di, j;
  long nnodes = 0;
  double logsum = 0, ssq = 0, cnt = (double)m->waic_count, lppd, pw;
  if (!m->waic_acc) return fail("online WAIC was not enabled");
  for (di = 0; di < m->t.n_decl; di++) {
    int n = m->t.decl[di].n;
    const double *acc = m->waic_acc[di];
    if (!acc) continue;
    for (j = 0; j < n; j++) { logsum += acc[j] + log(acc[n + j]); ssq += acc[3 * n + j]; }
    nnodes += n;
  }
  if (m->waic_count < 2) { out4[0] = out4[1] = out4[2] = NAN; out4[3] = cnt; return 0; }
  lppd = logsum - (double)nnodes * log(cnt); pw = ssq / (cnt - 1.0);
  out4[0] = -2.0 * (lppd - pw); out4[1] = lppd; out4[2] = pw; out4[3] = cnt;
  return 0;
}

/* initializeModel: see nimb200_initialize_model in include/nimb200.h (initializeModel.R:78-139) */
int ref_initialize_model(ref_model *m, const int32_t *decl_order, int32_t n_order, uint64_t seed) {
  int q, c, node;
  for (q = 0; q < n_order; q++) {
    int di = decl_order[q];
    const nimb200_decl *d;
    if (di < 0 || di >= m->t.n_decl) return fail("declaration index out of range");
    d = m->t.decl + di;
    if (d->value_kind != 1) continue;
    for (c = 0; c < m->C; c++) {
      double *st = m->state + (size_t)c * m->t.n_slots;
      uint32_t chain = (uint32_t)(m->first_chain + c);
      for (node = 0; node < d->n; node++) {
        double p0, p1, u, u2, nv;
        uint32_t sid = 0x7fff0000u + (uint32_t)di;
        if (!isnan(st[d->value_ref + node])) continue;
        p0 = get_param(m, st, di, node, 0);
        p1 = d->nparam > 1 ? get_param(m, st, di, node, 1) : 0.0;
        switch (d->dist) {
          case NIMB_DNORM: ref_normal2(seed, chain, 0, sid, node, 0, &u, &u2); nv = p0 + p1 * u; break;
          case NIMB_DGAMMA: nv = ref_rgamma(seed, chain, 0, sid, node, p0) * p1; break;
          case NIMB_DBETA: nv = ref_rbeta(seed, chain, 0, sid, node, p0, p1); break;
          case NIMB_DEXP: ref_unif2(seed, chain, 0, sid, node, 0, &u, &u2); nv = -ref_det_log(u) / p0; break;
          case NIMB_DLNORM: ref_normal2(seed, chain, 0, sid, node, 0, &u, &u2); nv = ref_det_exp(p0 + p1 * u); break;
          case NIMB_DLOGIS: ref_unif2(seed, chain, 0, sid, node, 0, &u, &u2); nv = p0 + p1 * (ref_det_log(u) - ref_det_log(1.0 - u)); break;
          case NIMB_DBINOM: nv = ref_rbinom(seed, chain, 0, sid, node, p0, p1); break;
          case NIMB_DPOIS: nv = ref_rpois(seed, chain, 0, sid, node, p0); break;
          default: return fail("initializeModel: a missing initial value cannot be simulated");
        }
        if (isnan(nv) && (d->dist == NIMB_DBINOM || d->dist == NIMB_DPOIS)) return fail("initializeModel: a missing initial value cannot be simulated");
        st[d->value_ref + node] = nv;
      }
    }
  }
  return 0;
}

int ref_draws(uint64_t seed, int32_t chain, int64_t iter, int32_t sampler, int32_t target, int32_t kind, double a,
              int32_t n, double *out) {
  int k;
  for (k = 0; k < n; k++) {
    double x0, x1;
    if (kind == 0) { ref_normal2(seed, chain, iter, sampler, target, k / 2, &x0, &x1); out[k] = (k & 1) ? x1 : x0; }
    else if (kind == 1) { ref_unif2(seed, chain, iter, sampler, target, k / 2, &x0, &x1); out[k] = (k & 1) ? x1 : x0; }
    else if (kind == 2) out[k] = ref_rgamma(seed, chain, iter, sampler, target + k, a);
    else if (kind == 3) out[k] = ref_rbeta(seed, chain, iter, sampler, target + k, a, 1.0 + 0.5 * a);
    else if (kind == 4) out[k] = ref_rpois(seed, chain, iter, sampler, target + k, a);
    else out[k] = ref_rbinom(seed, chain, iter, sampler, target + k, 25.0, a);
  }
  return 0;
}

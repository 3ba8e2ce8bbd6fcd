// engine.cu -- node-table interpreter kernels, samplers, MCMC driver and the C ABI (include/nimb200.h).
//
// Replaces, for the sampling path, the reference's per-node virtual dispatch
//   calculate / calculateDiff / getLogProb loops    P/inst/CppCode/accessorClasses.cpp:49-57, 237-264
//   nimCopy model <-> mvSaved                        P/inst/CppCode/accessorClasses.cpp:549-572
//   sampler_RW / RW_block / binary / conjugate       P/R/MCMC_samplers.R, P/R/MCMC_conjugacy.R
//   MCMC::run                                        P/R/MCMC_build.R:217-321
// with kernels that advance ALL chains at once over chain-major state.
//
// Two execution shapes:
//   sweep  one thread per (target, chain): the whole scalar sampler step, sequential over that target's
//          dependents (many conditionally independent targets, or few dependents);
//   wide   one target per chain with a large dependent set: (node tile x chain tile) blocks evaluate the
//          dependents, per-chain partial sums go to scratch and are reduced in fixed order, then a per-chain
//          kernel decides / commits.  Per-dependency block sums of the current state are cached so an MH
//          step costs one pass over the dependents, like calculateDiff against stored logProbs.
//
// Compiled with -fmad=false: a*b+c stays two roundings, as in the reference's x86-64 build.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <array>
#include <set>
#include "draws_dev.cuh"
#include "engine.cuh"
#include "nmath_dev.cuh"
#include "fastmath.cuh"

using namespace nimb;

// ------------------------------------------------------------------------------------------------ errors
static thread_local std::string g_err;
static int fail(const std::string& s) { g_err = s; return 1; }
namespace nimb { void set_error(const std::string& s) { g_err = s; } }  // for entry points in other translation units
#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      char b_[512];                                                                           \
      snprintf(b_, sizeof b_, "CUDA error %s at %s:%d", cudaGetErrorString(e_), __FILE__, __LINE__); \
      return fail(b_);                                                                        \
    }                                                                                         \
  } while (0)

// ------------------------------------------------------------------------------------------------ interpreter
__device__ __forceinline__ int idx_of(const DevTab& T, int mode, int arg, int i) {
  switch (mode) {
    case NIMB_IDX_I: return i;
    case NIMB_IDX_DIV: return i / arg;
    case NIMB_IDX_MOD: return i % arg;
    default: return T.iarr[arg][i];
  }
}

__device__ __forceinline__ double ld_state(const DevTab& T, int slot, int c, int ov_slot, double ov_val) {
  return slot == ov_slot ? ov_val : T.state[(size_t)slot * T.C + c];
}

__device__ double eval_prog(const DevTab& T, int c, int off, int len, int i, int j, int ov_slot, double ov_val) {
  const nimb200_op* ops = T.ops + off;
  if (len >= 1 && len <= 2) {  // a constant or a chain variable, possibly inverted (rate -> scale): no stack, no dispatch loop
    const nimb200_op o0 = ops[0];
    if (o0.op == NIMB_OP_CONST || o0.op == NIMB_OP_CVAR) {
      const double v = o0.op == NIMB_OP_CONST ? o0.k : ld_state(T, o0.a, c, ov_slot, ov_val);
      if (len == 1) return v;
      if (ops[1].op == NIMB_OP_INV) return 1. / v;
    }
  }
  double stk[16];
  int sp = 0;
  for (int q = 0; q < len; q++) {
    const nimb200_op o = ops[q];
    switch (o.op) {
      case NIMB_OP_CONST: stk[sp++] = o.k; break;
      case NIMB_OP_DATA: stk[sp++] = T.darr[o.a][idx_of(T, o.b, o.c, i)]; break;
      case NIMB_OP_CVAR: stk[sp++] = ld_state(T, o.a, c, ov_slot, ov_val); break;
      case NIMB_OP_CVAR_IDX: {
        double v = stk[sp - 1];
        stk[sp - 1] = (v >= 1.0 && v <= (double)o.b && v == floor(v)) ? ld_state(T, o.a + (int)v - 1, c, ov_slot, ov_val) : NAN;
        break;
      }
      case NIMB_OP_DATA_IDX: {
        double v = stk[sp - 1];
        stk[sp - 1] = (v >= 1.0 && v <= (double)o.b && v == floor(v)) ? T.darr[o.a][(int)v - 1] : NAN;
        break;
      }
      case NIMB_OP_CVARN: stk[sp++] = ld_state(T, o.a + idx_of(T, o.b, o.c, i), c, ov_slot, ov_val); break;
      case NIMB_OP_INPROD: {
        const double* x = T.darr[o.a] + (size_t)i * o.c;
        int p = (int)o.k;
        double s = 0.0;
        for (int jj = 0; jj < p; jj++) s += x[jj] * ld_state(T, o.b + jj, c, ov_slot, ov_val);
        stk[sp++] = s;
        break;
      }
      case NIMB_OP_DATAJ: stk[sp++] = T.darr[o.a][(size_t)i * o.c + j]; break;
      case NIMB_OP_CVARJ: stk[sp++] = ld_state(T, o.a + j, c, ov_slot, ov_val); break;
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

__device__ __forceinline__ double node_value(const DevTab& T, const nimb200_decl& d, int i, int c, int ov_slot,
                                             double ov_val) {
  return d.value_kind ? ld_state(T, d.value_ref + i, c, ov_slot, ov_val) : T.darr[d.value_ref][i];
}

// getParam_0D_double (accessorClasses.h:210-224): canonical parameter or one of the derived alternatives
__device__ double get_param(const DevTab& T, int di, int i, int c, int code, int ov_slot, double ov_val) {
  const nimb200_decl& d = T.decl[di];
  if (code < NIMB200_MAX_PARAM) return eval_prog(T, c, d.prog_off[code], d.prog_len[code], i, 0, ov_slot, ov_val);
  double p1 = eval_prog(T, c, d.prog_off[1], d.prog_len[1], i, 0, ov_slot, ov_val);
  if (code == NIMB_PARAM_RATE) return 1.0 / p1;
  if (code == NIMB_PARAM_TAU) return pow(p1, -2.0);
  return NAN;
}

// one nodeFunction::calculate body: d<dist>(value, params, log = 1)
__device__ double calc_node(const DevTab& T, int di, int i, int c, int ov_slot, double ov_val) {
  const nimb200_decl& d = T.decl[di];
  double x = node_value(T, d, i, c, ov_slot, ov_val);
  if (d.dist == NIMB_DCAT) {  // P/inst/CppCode/dists.cpp:914-934
    int K = d.fam_arg[0];
    double sum = 0.0, sel = 0.0;
    bool anynan = isnan(x), anyneg = false;
    int xi = isnan(x) ? 0 : (int)nearbyint(x);
    for (int j = 0; j < K; j++) {
      double pj = eval_prog(T, c, d.prog_off[0], d.prog_len[0], i, j, ov_slot, ov_val);
      anynan |= isnan(pj);
      anyneg |= pj < 0;
      sum += pj;
      if (j == xi - 1) sel = pj;
    }
    if (anynan || anyneg) return NAN;
    if (r_nonint(x)) return -INFINITY;
    x = nearbyint(x);
    if (x > K || x < 1) return -INFINITY;
    return log(sel) - log(sum);
  }
  double p0 = eval_prog(T, c, d.prog_off[0], d.prog_len[0], i, 0, ov_slot, ov_val), p1 = 0;
  if (d.nparam > 1) p1 = eval_prog(T, c, d.prog_off[1], d.prog_len[1], i, 0, ov_slot, ov_val);
  switch (d.dist) {
    case NIMB_DNORM: return dnorm(x, p0, p1);
    case NIMB_DGAMMA: return dgamma(x, p0, p1);
    case NIMB_DPOIS: return dpois(x, p0);
    case NIMB_DBINOM: return dbinom(x, p0, p1);
    case NIMB_DBETA: return dbeta(x, p0, p1);
    case NIMB_DLOGIS: return dlogis(x, p0, p1);
    case NIMB_DEXP: return dexp_nimble(x, p0);
    case NIMB_DLNORM: return dlnorm(x, p0, p1);
  }
  return NAN;
}

// Parameters whose programs do not read anything indexed by the node (sd = sig_e, shape = alpha, ...) are evaluated
// once per thread instead of once per node; for dnorm the log(sd) term is carried along.  Results are bit-identical to
// calc_node(): the same operations on the same values, only not repeated.
struct Hoist { bool h0, h1, gfast; double p0, p1, l1; };  // l1: log(sd) of dnorm, or -shape log(scale) - lgamma(shape) of dgamma
__device__ __forceinline__ Hoist make_hoist(const DevTab& T, int di, int c, int ov_slot, double ov_val) {
  const nimb200_decl& d = T.decl[di];
  Hoist H; H.p0 = H.p1 = H.l1 = 0.0; H.gfast = false;
  H.h0 = d.dist != NIMB_DCAT && T.pconst[3 * di] != 0;
  H.h1 = d.dist != NIMB_DCAT && d.nparam > 1 && T.pconst[3 * di + 1] != 0;
  if (H.h0) H.p0 = eval_prog(T, c, d.prog_off[0], d.prog_len[0], 0, 0, ov_slot, ov_val);
  if (H.h1) {
    H.p1 = eval_prog(T, c, d.prog_off[1], d.prog_len[1], 0, 0, ov_slot, ov_val);
    if (d.dist == NIMB_DNORM) H.l1 = log(H.p1);
    // dgamma with chain-level shape and scale: the node-independent part once per thread, then (shape - 1) log x - x / scale per
    // node instead of nmath's dpois_raw route (same density; the closed form loses digits to cancellation only for large
    // shapes, which keep the exact route)
    if (d.dist == NIMB_DGAMMA && H.h0 && H.p0 > 0.0 && H.p0 <= 64.0 && H.p1 > 0.0 && isfinite(H.p1)) {
      H.gfast = true;
      H.l1 = -H.p0 * log(H.p1) - lgamma(H.p0);
    }
  }
  return H;
}
__device__ __forceinline__ double dnorm_logsd(double x, double mu, double sigma, double logsigma) {  // dnorm() with log(sigma) given
  if (isnan(x) || isnan(mu) || isnan(sigma)) return x + mu + sigma;
  if (sigma < 0) return NAN;
  if (!isfinite(sigma)) return -INFINITY;
  if (!isfinite(x) && mu == x) return NAN;
  if (sigma == 0) return (x == mu) ? INFINITY : -INFINITY;
  x = (x - mu) / sigma;
  if (!isfinite(x)) return -INFINITY;
  x = fabs(x);
  if (x >= 2 * sqrt(DBL_MAX)) return -INFINITY;
  return -(NIMB_LN_SQRT_2PI + 0.5 * x * x + logsigma);
}
__device__ double calc_node_h(const DevTab& T, int di, int i, int c, int ov_slot, double ov_val, const Hoist& H) {
  const nimb200_decl& d = T.decl[di];
  if (d.dist == NIMB_DCAT) return calc_node(T, di, i, c, ov_slot, ov_val);
  double x = node_value(T, d, i, c, ov_slot, ov_val);
  double p0 = H.h0 ? H.p0 : eval_prog(T, c, d.prog_off[0], d.prog_len[0], i, 0, ov_slot, ov_val), p1 = 0;
  if (d.nparam > 1) p1 = H.h1 ? H.p1 : eval_prog(T, c, d.prog_off[1], d.prog_len[1], i, 0, ov_slot, ov_val);
  switch (d.dist) {
    case NIMB_DNORM: return H.h1 ? dnorm_logsd(x, p0, p1, H.l1) : dnorm(x, p0, p1);
    case NIMB_DGAMMA: return (H.gfast && x > 0.0 && isfinite(x)) ? fma(p0 - 1.0, log(x), H.l1) - x / p1 : dgamma(x, p0, p1);
    case NIMB_DPOIS: return dpois(x, p0);
    case NIMB_DBINOM: return dbinom(x, p0, p1);
    case NIMB_DBETA: return dbeta(x, p0, p1);
    case NIMB_DLOGIS: return dlogis(x, p0, p1);
    case NIMB_DEXP: return dexp_nimble(x, p0);
    case NIMB_DLNORM: return dlnorm(x, p0, p1);
  }
  return NAN;
}

// thread index -> (target, chain), chains fastest.  A 64-bit division costs ~60 instructions; almost every launch has fewer than
// 2^32 threads, where one 32-bit division does (warp-uniform choice)
__device__ __forceinline__ void split_index(size_t gid, int C, size_t& tq, int& c) {
  if ((gid >> 32) == 0) {
    const unsigned g = (unsigned)gid, q = g / (unsigned)C;
    tq = q;
    c = (int)(g - q * (unsigned)C);
  } else {
    tq = gid / (size_t)C;
    c = (int)(gid - tq * (size_t)C);
  }
}

struct Sel { int lo, hi; const int32_t* idx; int nlo, nhi; };  // [nlo,nhi): node instances this rank owns (data sharding)
__device__ __forceinline__ Sel sel_of(const DevTab& T, const nimb200_dep& dp, int t) {
  Sel s; s.idx = nullptr;
  switch (dp.sel) {
    case NIMB_SEL_ALL: s.lo = 0; s.hi = T.decl[dp.decl].n; break;
    case NIMB_SEL_SAME: s.lo = dp.a + t; s.hi = s.lo + 1; break;
    case NIMB_SEL_STRIDE: s.lo = t * dp.a; s.hi = s.lo + dp.a; break;
    case NIMB_SEL_SEG: s.lo = T.iarr[dp.a][t]; s.hi = T.iarr[dp.a][t + 1]; break;
    case NIMB_SEL_CSR: s.lo = T.iarr[dp.a][t]; s.hi = T.iarr[dp.a][t + 1]; s.idx = T.iarr[dp.b]; break;
    default: s.lo = dp.a; s.hi = dp.b; break;
  }
  s.nlo = T.shard_lo[dp.decl]; s.nhi = T.shard_hi[dp.decl];
  if (!s.idx) { s.lo = max(s.lo, s.nlo); s.hi = min(s.hi, s.nhi); }
  return s;
}
// does this rank account for declaration `decl` in a quantity that is summed over ranks?  Sharded declarations:
// every rank, its own slice.  Replicated declarations: rank 0 only.
__device__ __forceinline__ bool counts_here(const DevTab& T, int decl) { return T.rank == 0 || T.sharded[decl]; }

// conjugate contribution of one dependent node (P/R/MCMC_conjugacy.R:80-93, 109-115; probing :766-820)
// the dependent's parameter that the posterior needs besides the linked one (tau | mean of dnorm, shape of dgamma);
// *hoisted is set when its program does not depend on the node, so the caller may reuse the value
// conjugate kind: 0 = dnorm prior, 1 = dgamma prior, 2 = dbeta prior
__host__ __device__ __forceinline__ int conj_kind(int sampler_type) {
  return sampler_type == NIMB_SAMPLER_CONJ_GAMMA ? 1 : (sampler_type == NIMB_SAMPLER_CONJ_BETA ? 2 : 0);
}
__device__ __forceinline__ double conj_par(const DevTab& T, const nimb200_dep& dp, int node, int c, int gam, bool* hoistable) {
  const nimb200_decl& dd = T.decl[dp.decl];
  if (dd.dist == NIMB_DBINOM) { *hoistable = T.pconst[3 * dp.decl] != 0; return get_param(T, dp.decl, node, c, 0, -1, 0.0); }  // size
  if (dd.dist == NIMB_DNORM || dd.dist == NIMB_DLNORM) {
    *hoistable = T.pconst[3 * dp.decl + (gam == 1 ? 0 : 1)] != 0;
    return gam == 1 ? get_param(T, dp.decl, node, c, 0, -1, 0.0) : get_param(T, dp.decl, node, c, NIMB_PARAM_TAU, -1, 0.0);
  }
  if (dd.dist == NIMB_DGAMMA) { *hoistable = T.pconst[3 * dp.decl] != 0; return get_param(T, dp.decl, node, c, 0, -1, 0.0); }
  *hoistable = true;
  return 0.0;
}
__device__ __forceinline__ void conj_contrib(const DevTab& T, const nimb200_dep& dp, int node, int c, int slot, int gam,
                                             double& c0, double& c1, bool have_par = false, double par_in = 0.0) {
  const nimb200_decl& dd = T.decl[dp.decl];
  double value = node_value(T, dd, node, c, -1, 0.0);
  double par = 0, offset = 0, coeff = 1;
  bool dummy;
  par = have_par ? par_in : conj_par(T, dp, node, c, gam, &dummy);
  if (dp.link == NIMB_LINK_ADDITIVE || dp.link == NIMB_LINK_LINEAR) offset = get_param(T, dp.decl, node, c, dp.param, slot, 0.0);
  if (dp.link == NIMB_LINK_LINEAR) coeff = get_param(T, dp.decl, node, c, dp.param, slot, 1.0) - offset;
  if (dp.link == NIMB_LINK_MULTIPLICATIVE) coeff = get_param(T, dp.decl, node, c, dp.param, slot, 1.0);
  if (gam == 2) {  // MCMC_conjugacy.R:4-8: dbern (value, 1 - value), dbin (value, size - value)
    c0 += value; c1 += par - value;
  } else if (gam == 1) {
    if (dd.dist == NIMB_DPOIS) { c0 += value; c1 += coeff; }
    else if (dd.dist == NIMB_DGAMMA) { c0 += par; c1 += coeff * value; }
    else if (dd.dist == NIMB_DEXP) { c0 += 1.0; c1 += coeff * value; }
    else if (dd.dist == NIMB_DNORM) { c0 += 1. / 2; c1 += coeff / 2 * ((value - par) * (value - par)); }
    else if (dd.dist == NIMB_DLNORM) { c0 += 1. / 2; c1 += coeff / 2 * ((log(value) - par) * (log(value) - par)); }
  } else {
    if (dd.dist == NIMB_DLNORM) value = log(value);   // MCMC_conjugacy.R:113: a dlnorm dependent contributes through log(value)
    c0 += coeff * (value - offset) * par;
    c1 += coeff * coeff * par;
  }
}

__device__ __forceinline__ DrawKey make_key(const DevTab& T, uint64_t seed, uint64_t iter, int sid, int t, int c) {
  DrawKey k; k.seed = seed; k.chain = (uint32_t)(T.first_chain + c); k.iter = T.ctr ? (uint64_t)T.ctr[0] : iter;
  k.sampler = (uint32_t)sid; k.target = (uint32_t)t;
  return k;
}

// decide(): P/inst/CppCode/Utils.cpp:54-63 (the uniform is only consulted when it is needed)
__device__ __forceinline__ bool decide(double lMHr, const DrawKey& k, uint32_t call) {
  if (isnan(lMHr)) return false;
  if (lMHr > 0) return true;
  double u, u2;
  unif2(k, call, u, u2);
  return u < exp(lMHr);
}

// sampler_RW::adaptiveProcedure, P/R/MCMC_samplers.R:388-420
__device__ __forceinline__ void rw_adapt(const nimb200_sampler& sp, const SamplerDev& S, size_t q, bool jump, int gridDimC) {
  int ran = S.timesRan[q] + 1, acc = S.timesAccepted[q] + (jump ? 1 : 0);
  if (ran % sp.adapt_interval == 0) {
    double acceptanceRate = (double)acc / (double)ran;
    int adapted = S.timesAdapted[q] + 1;
    S.timesAdapted[q] = adapted;
    if (S.hist && adapted <= S.hist_cap) {  // scaleHistory / acceptanceHistory, MCMC_samplers.R:394-399
      const size_t TC = (size_t)sp.n_targets * gridDimC;
      S.hist[((size_t)(adapted - 1) * S.hist_K + 0) * TC + q] = S.scale[q];
      S.hist[((size_t)(adapted - 1) * S.hist_K + 1) * TC + q] = acceptanceRate;
    }
    double gamma1 = 1 / pow((double)adapted + 3, sp.adapt_factor_exponent);
    double gamma2 = 10 * gamma1;
    double adaptFactor = exp(gamma2 * (acceptanceRate - 0.44));
    double sc = S.scale[q] * adaptFactor;
    if (sp.reflective && sc >= 0.5 * (sp.upper - sp.lower)) sc = 0.5 * (sp.upper - sp.lower);
    S.scale[q] = sc;
    ran = 0; acc = 0;
  }
  S.timesRan[q] = ran;
  S.timesAccepted[q] = acc;
}

__device__ __forceinline__ double rw_proposal(const nimb200_sampler& sp, double cur, double scale, double z, double& propLogScale) {
  double prop;
  propLogScale = 0;
  if (sp.log_scale) {
    propLogScale = 0.0 + scale * z;
    prop = cur * exp(propLogScale);
  } else
    prop = cur + scale * z;
  if (sp.reflective) {
    while (prop < sp.lower || prop > sp.upper) {
      if (prop < sp.lower) prop = 2 * sp.lower - prop;
      if (prop > sp.upper) prop = 2 * sp.upper - prop;
    }
  }
  return prop;
}

#include "fused_pois.cuh"
#include "fused_norm.cuh"
#include "fused_bern.cuh"
#include "peer_xchg.cuh"

// ------------------------------------------------------------------------------------------------ sweep kernels
// One thread = one (target t, chain c): the complete scalar sampler step.  Chains are the fast index, so the
// 32 lanes of a warp read the same data node and 32 consecutive chain copies of each state slot.
// PHASE 0: everything in one kernel.  Data-sharded models split it: PHASE 1 accumulates this rank's share of the
// per-(target, chain) sums into buf[t][k][c] (k < 3), the host all-reduces buf, PHASE 2 decides / draws from it.
// LANES = 32: one WARP per (target, chain), its lanes stride over the dependent nodes and the partial sums meet in a fixed-order
// butterfly -- used when there are few (target, chain) pairs but each has many dependents (tau_b / mu_b over 50 betas).
// gthread: index of the calling thread among the (target, chain, lane) work items of the sampler
template <int PHASE, int LANES>
__device__ __forceinline__ void sweep_body(const DevTab& T, int sid, uint64_t seed, uint64_t iter, const SamplerDev& S, double* buf,
                                           size_t gthread) {
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = gthread / LANES;
  const int lane = LANES == 1 ? 0 : (int)(gthread & (LANES - 1));
  int c;
  size_t tq;
  split_index(gid, T.C, tq, c);
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq;
  int slot = sp.target_slot + t;
  size_t q = (size_t)t * T.C + c;
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double* sv = T.state + (size_t)slot * T.C + c;
  double cur = *sv;
  double v0 = 0, v1 = 0, v2 = 0;
  double* bp = buf + ((size_t)t * 3) * T.C + c;
  if (PHASE == 2) { v0 = bp[0]; v1 = bp[(size_t)T.C]; v2 = bp[2 * (size_t)T.C]; }

  if (sp.type == NIMB_SAMPLER_POSTERIOR_PREDICTIVE) {  // MCMC_samplers.R:97-101: simulate the target, nothing to decide (its network follows as further table samplers)
    if (PHASE == 1 || lane != 0) return;
    const nimb200_dep tp = T.deps[sp.dep_off];
    const int tnode = (tp.sel == NIMB_SEL_SAME ? tp.a + t : tp.a);
    const nimb200_decl& d = T.decl[tp.decl];
    const double p0 = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
    const double p1 = d.nparam > 1 ? get_param(T, tp.decl, tnode, c, 1, -1, 0.0) : 0.0;
    double u, u2, nv = NAN;
    switch (d.dist) {
      case NIMB_DNORM: normal2(key, 0, u, u2); nv = p0 + p1 * u; break;
      case NIMB_DGAMMA: nv = rgamma1(key, p0) * p1; break;
      case NIMB_DBETA: nv = rbeta1(key, p0, p1); break;
      case NIMB_DEXP: unif2(key, 0, u, u2); nv = -det_log(u) / p0; break;
      case NIMB_DLNORM: normal2(key, 0, u, u2); nv = det_exp(p0 + p1 * u); break;
      case NIMB_DLOGIS: unif2(key, 0, u, u2); nv = p0 + p1 * (det_log(u) - det_log(1.0 - u)); break;
      case NIMB_DBINOM: nv = rbinom_inv(key, p0, p1); break;
      case NIMB_DPOIS: nv = rpois_inv(key, p0); break;
    }
    *sv = nv;
    return;
  }
  if (sp.type == NIMB_SAMPLER_RW) {  // MCMC_samplers.R:352-386
    double z, z1, propLogScale;
    normal2(key, 0, z, z1);
    double prop = rw_proposal(sp, cur, S.scale[q], z, propLogScale);
    if (PHASE != 2) {
      for (int e = 0; e < sp.n_dep; e++) {
        const nimb200_dep dp = T.deps[sp.dep_off + e];
        if (PHASE == 1 && !counts_here(T, dp.decl)) continue;
        Sel s = sel_of(T, dp, t);
        Hoist Hn = make_hoist(T, dp.decl, c, slot, prop), Ho = make_hoist(T, dp.decl, c, -1, 0.0);
        for (int k = s.lo + lane; k < s.hi; k += LANES) {
          int node = s.idx ? s.idx[k] : k;
          if (s.idx && (node < s.nlo || node >= s.nhi)) continue;
          double nw = calc_node_h(T, dp.decl, node, c, slot, prop, Hn);
          double od = calc_node_h(T, dp.decl, node, c, -1, 0.0, Ho);
          if (dp.stage == 0) v0 += nw - od; else v1 += nw - od;
        }
      }
    }
    if (LANES > 1) {
#pragma unroll
      for (int o = LANES / 2; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        v2 += __shfl_xor_sync(0xffffffffu, v2, o);
      }
      if (lane != 0) return;
    }
    if (PHASE == 1) { bp[0] = v0; bp[(size_t)T.C] = v1; bp[2 * (size_t)T.C] = 0.0; return; }
    bool jump;
    if (v0 == -INFINITY) jump = false;
    else jump = decide(v0 + v1 + propLogScale, key, 1);
    if (jump) *sv = prop;
    if (sp.adaptive) rw_adapt(sp, S, q, jump, T.C);
    return;
  }
  if (sp.type == NIMB_SAMPLER_BINARY) {  // MCMC_samplers.R:127-151
    double other = 1 - cur;
    if (PHASE != 2) {
      for (int e = 0; e < sp.n_dep; e++) {
        const nimb200_dep dp = T.deps[sp.dep_off + e];
        if (PHASE == 1 && !counts_here(T, dp.decl)) continue;
        Sel s = sel_of(T, dp, t);
        Hoist Hn = make_hoist(T, dp.decl, c, slot, other), Ho = make_hoist(T, dp.decl, c, -1, 0.0);
        for (int k = s.lo + lane; k < s.hi; k += LANES) {
          int node = s.idx ? s.idx[k] : k;
          if (s.idx && (node < s.nlo || node >= s.nhi)) continue;
          v0 += calc_node_h(T, dp.decl, node, c, -1, 0.0, Ho);
          double o = calc_node_h(T, dp.decl, node, c, slot, other, Hn);
          if (dp.stage == 0) v1 += o; else v2 += o;
        }
      }
    }
    if (LANES > 1) {
#pragma unroll
      for (int o = LANES / 2; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        v2 += __shfl_xor_sync(0xffffffffu, v2, o);
      }
      if (lane != 0) return;
    }
    if (PHASE == 1) { bp[0] = v0; bp[(size_t)T.C] = v1; bp[2 * (size_t)T.C] = v2; return; }
    double currentLogProb = v0, otherPrior = v1, otherDeps = v2;
    double otherLogProb = (otherPrior == -INFINITY) ? otherPrior : otherPrior + otherDeps;
    double acceptanceProb = 1 / (exp(currentLogProb - otherLogProb) + 1);
    double u, u2;
    unif2(key, 0, u, u2);
    bool jump = (!isnan(acceptanceProb)) & (u < acceptanceProb);
    if (jump) *sv = other;
    return;
  }
  // conjugate: MCMC_conjugacy.R:651-692
  const int gam = conj_kind(sp.type);
  if (PHASE != 2) {
    for (int e = 0; e < sp.n_dep; e++) {
      const nimb200_dep dp = T.deps[sp.dep_off + e];
      if (dp.stage != 1) continue;
      if (PHASE == 1 && !counts_here(T, dp.decl)) continue;
      Sel s = sel_of(T, dp, t);
      bool hp = false;
      double par = conj_par(T, dp, 0, c, gam, &hp);
      for (int k = s.lo + lane; k < s.hi; k += LANES) {
        int node = s.idx ? s.idx[k] : k;
        if (s.idx && (node < s.nlo || node >= s.nhi)) continue;
        conj_contrib(T, dp, node, c, slot, gam, v0, v1, hp, par);
      }
    }
  }
  if (LANES > 1) {
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) {
      v0 += __shfl_xor_sync(0xffffffffu, v0, o);
      v1 += __shfl_xor_sync(0xffffffffu, v1, o);
      v2 += __shfl_xor_sync(0xffffffffu, v2, o);
    }
    if (lane != 0) return;
  }
  if (PHASE == 1) { bp[0] = v0; bp[(size_t)T.C] = v1; bp[2 * (size_t)T.C] = 0.0; return; }
  const nimb200_dep tp = T.deps[sp.dep_off];
  int tnode = (tp.sel == NIMB_SEL_SAME ? tp.a + t : tp.a);
  double prior0 = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
  double prior1 = get_param(T, tp.decl, tnode, c, gam == 2 ? 1 : (gam == 1 ? NIMB_PARAM_RATE : NIMB_PARAM_TAU), -1, 0.0);
  double nv;
  if (gam == 2) {
    nv = rbeta1(key, prior0 + v0, prior1 + v1);
  } else if (gam == 1) {
    double shape = prior0 + v0, scale = 1 / (prior1 + v1);
    nv = rgamma1(key, shape) * scale;
  } else {
    double mean = (prior0 * prior1 + v0) / (prior1 + v1), sd = pow(prior1 + v1, -0.5), z, z1;
    normal2(key, 0, z, z1);
    nv = mean + sd * z;
  }
  *sv = nv;
}

template <int PHASE, int LANES>
__global__ void k_sweep(DevTab T, int sid, uint64_t seed, uint64_t iter, SamplerDev S, double* buf) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  sweep_body<PHASE, LANES>(T, sid, seed, iter, S, buf, (size_t)blockIdx.x * blockDim.x + threadIdx.x);
}

// Small models (BASELINE config 1, the BUGS 'pump' example: 12 latent nodes, a handful of chains): an iteration is three or four
// launches of a few dozen threads each and costs nothing but launch latency (~10 us per kernel even from a replayed graph).
// When every sampler of the model is a scalar sweep with little work, ONE block runs whole iterations -- sampler after sampler
// with a block barrier in between, then the monitor row and the iteration counter -- and n_it of them per launch.  Same
// sweep_body, same draws, same results as the launch-per-sampler path.
struct SmallPlan { int n; int sid[24]; int lanes[24]; };
__global__ void __launch_bounds__(1024)
k_small_iters(DevTab T, SmallPlan P, const SamplerDev* __restrict__ sdev, uint64_t seed, int n_it, const int32_t* __restrict__ mon,
              int nmon, const long long* __restrict__ rp /* RunParams as words */, double* samples, unsigned long long* ctr) {
  const long long rows_before = rp[0], iter0 = rp[1], nburnin = rp[2];
  const int thin = (int)(rp[3] & 0xffffffffll);
  for (int it = 0; it < n_it; it++) {
    for (int j = 0; j < P.n; j++) {
      const int sid = P.sid[j];
      const size_t total = (size_t)T.samplers[sid].n_targets * T.C * P.lanes[j];
      const SamplerDev S = sdev[sid];
      if (P.lanes[j] == 32) { for (size_t g = threadIdx.x; g < total; g += blockDim.x) sweep_body<0, 32>(T, sid, seed, 0, S, nullptr, g); }
      else { for (size_t g = threadIdx.x; g < total; g += blockDim.x) sweep_body<0, 1>(T, sid, seed, 0, S, nullptr, g); }
      __syncthreads();
    }
    const long long itl = (long long)ctr[0] - iter0;
    if (itl > nburnin && (itl - nburnin) % thin == 0) {
      const long long row = rows_before + (itl - nburnin) / thin - 1;
      for (size_t g = threadIdx.x; g < (size_t)nmon * T.C; g += blockDim.x)
        samples[(size_t)row * nmon * T.C + g] = T.state[(size_t)mon[g / T.C] * T.C + (g % T.C)];
    }
    __syncthreads();
    if (threadIdx.x == 0) ctr[0] += 1;
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------ slice sampler
// sampler_slice (MCMC_samplers.R:864-972), one thread per (target, chain): the number of log-density evaluations is data
// dependent (stepping out, then shrinkage), so the whole update is one loop in one thread over the target's dependents with the
// candidate value in the override register.  Uniforms: call 0 = (auxiliary exponential, interval offset), call 1 = (split of
// maxSteps, first candidate), call 2 + j = j-th shrinkage candidate.
__device__ double slice_lp_stage(const DevTab& T, const nimb200_sampler& sp, int t, int c, int slot, double value, int stage) {
  double lp = 0.0;
  for (int e = 0; e < sp.n_dep; e++) {
    const nimb200_dep dp = T.deps[sp.dep_off + e];
    if (dp.stage != stage) continue;
    Sel s = sel_of(T, dp, t);
    for (int k = s.lo; k < s.hi; k++) {
      int node = s.idx ? s.idx[k] : k;
      lp += calc_node(T, dp.decl, node, c, slot, value);
    }
  }
  return lp;
}
__device__ double slice_lp(const DevTab& T, const nimb200_sampler& sp, int t, int c, int slot, double value) {
  const double lp0 = slice_lp_stage(T, sp, t, c, slot, value, 0);
  if (lp0 == -INFINITY) return -INFINITY;  // setAndCalculateTarget: prior -Inf short-circuits
  return lp0 + slice_lp_stage(T, sp, t, c, slot, value, 1);
}
__global__ void k_slice(DevTab T, int sid, uint64_t seed, uint64_t iter, SamplerDev S) {
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c;
  size_t tq;
  split_index(gid, T.C, tq, c);
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq, slot = sp.target_slot + t;
  size_t q = (size_t)t * T.C + c;
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double* sv = T.state + (size_t)slot * T.C + c;
  const int tdist = T.decl[T.deps[sp.dep_off].decl].dist;
  const bool discrete = tdist == NIMB_DPOIS || tdist == NIMB_DBINOM || tdist == NIMB_DCAT;
  const double eps = 1e-15, x0 = *sv, width = S.scale[q];
  auto f = [&](double v) { return slice_lp(T, sp, t, c, slot, discrete ? floor(v) : v); };
  double ua, ub, uc, ud;
  unif2(key, 0, ua, ub);
  unif2(key, 1, uc, ud);
  const double u = f(x0) - (-det_log(ua));
  double L = x0 - ub * width, R = L + width;
  int maxStepsL = (int)floor(uc * (double)sp.tries), maxStepsR = sp.tries - 1 - maxStepsL;
  double lp = f(L);
  while (maxStepsL > 0 && !isnan(lp) && lp >= u) { L = L - width; lp = f(L); maxStepsL--; }
  lp = f(R);
  while (maxStepsR > 0 && !isnan(lp) && lp >= u) { R = R + width; lp = f(R); maxStepsR--; }
  double x1 = L + ud * (R - L);
  lp = f(x1);
  int nc = 0;
  while ((isnan(lp) || lp < u) && (R - L) / (fabs(R) + fabs(L) + eps) > eps && nc < sp.pad) {
    if (x1 < x0) L = x1; else R = x1;
    double ue, uf;
    unif2(key, 2 + nc, ue, uf);
    x1 = L + ue * (R - L);
    lp = f(x1);
    nc++;
  }
  if ((R - L) / (fabs(R) + fabs(L) + eps) <= eps || nc == sp.pad) return;  // invalid state / collapsed interval: keep the old value
  *sv = discrete ? floor(x1) : x1;
  if (sp.adaptive) {  // adaptiveProcedure: width -> 2 x mean jump, with weight (3/4)^timesAdapted
    int ran = S.timesRan[q] + 1;
    double sum = S.propCov[q] + fabs(x1 - x0);
    if (ran % sp.adapt_interval == 0) {
      double af = 1.0;
      for (int i = 0; i < S.timesAdapted[q]; i++) af = af * 0.75;
      double meanJump = sum / (double)ran;
      S.scale[q] = width + (2 * meanJump - width) * af;
      S.timesAdapted[q] += 1;
      ran = 0; sum = 0.0;
    }
    S.timesRan[q] = ran;
    S.propCov[q] = sum;
  }
}

// sampler_categorical (MCMC_samplers.R:187-216), one thread per (target, chain): logProbs[i] for every category i (the current
// one included: the state is up to date, so recomputing equals getLogProb), -Inf for a -Inf or NaN prior or a NaN total, then
// probs = exp(logProbs - max) and rcat's linear scan on u * sum(probs) (dists.cpp:953-959).  exp is the deterministic one of the
// draw stream so that the scan compares the same bits on both sides.
__global__ void k_categorical(DevTab T, int sid, uint64_t seed, uint64_t iter) {
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c;
  size_t tq;
  split_index(gid, T.C, tq, c);
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq, slot = sp.target_slot + t;
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double* sv = T.state + (size_t)slot * T.C + c;
  const nimb200_dep d0 = T.deps[sp.dep_off];
  const nimb200_decl& td = T.decl[d0.decl];
  const int K = td.fam_arg[0];
  double lps[NIMB200_MAX_CATEGORIES];
  double maxLP = -INFINITY;
  // The target's own dcat density is log(p_i) - log(sum p) for category i: the K probabilities and their sum do not depend on
  // the candidate, so they are evaluated once instead of once per category (the same operations in the same order as
  // calc_node's, hence the same bits) when the prior stage is exactly the target's own node.
  bool own_prior = d0.stage == 0 && d0.sel == NIMB_SEL_SAME && td.dist == NIMB_DCAT;
  for (int e = 1; e < sp.n_dep && own_prior; e++) own_prior = T.deps[sp.dep_off + e].stage != 0;
  double logsum = 0.0;
  bool bad = false;
  if (own_prior) {
    const int tnode = d0.a + t;
    double sum = 0.0;
    for (int j = 0; j < K; j++) {
      const double pj = eval_prog(T, c, td.prog_off[0], td.prog_len[0], tnode, j, -1, 0.0);
      bad |= isnan(pj) || pj < 0;
      sum += pj;
      lps[j] = pj;
    }
    logsum = log(sum);
  }
  if (own_prior) {
    // dependents: entry by entry and node by node on the outside (selection and entry loaded once), categories on the inside;
    // every category still adds its dependents up in the same order as slice_lp does
    double dsum[NIMB200_MAX_CATEGORIES];
    for (int i = 0; i < K; i++) dsum[i] = 0.0;
    for (int e = 1; e < sp.n_dep; e++) {
      const nimb200_dep dp = T.deps[sp.dep_off + e];
      const Sel s = sel_of(T, dp, t);
      for (int k = s.lo; k < s.hi; k++) {
        const int node = s.idx ? s.idx[k] : k;
        for (int i = 0; i < K; i++) dsum[i] += calc_node(T, dp.decl, node, c, slot, (double)(i + 1));
      }
    }
    for (int i = 0; i < K; i++) {
      const double lp0 = 0.0 + (bad ? NAN : log(lps[i]) - logsum);
      double lp = lp0 == -INFINITY ? -INFINITY : lp0 + dsum[i];   // setAndCalculateTarget: a -Inf prior short-circuits
      if (isnan(lp)) lp = -INFINITY;
      lps[i] = lp;
      maxLP = fmax(maxLP, lp);
    }
  } else {
    for (int i = 1; i <= K; i++) {
      double lp = slice_lp(T, sp, t, c, slot, (double)i);  // prior first (-Inf short-circuits), then the dependents
      if (isnan(lp)) lp = -INFINITY;
      lps[i - 1] = lp;
      maxLP = fmax(maxLP, lp);
    }
  }
  double sum = 0.0;
  for (int i = 0; i < K; i++) { lps[i] = det_exp(lps[i] - maxLP); sum += lps[i]; }
  if (isnan(sum)) return;  // rcat gives NaN: newValue is NA, the state stays (MCMC_samplers.R:205)
  double u, u2;
  unif2(key, 0, u, u2);
  u = u * sum;
  double cum = lps[0];
  int value = 1;
  while (u > cum && value < K) { cum += lps[value]; value++; }
  *sv = (double)value;
}

// initializeModel: simulate missing (NaN) values of one latent declaration from its own distribution (see nimb200.h)
__device__ __forceinline__ bool simulate_node(const DevTab& T, int di, int node, int c, const DrawKey& key, double& nv) {
  const nimb200_decl& d = T.decl[di];
  const double p0 = get_param(T, di, node, c, 0, -1, 0.0);
  const double p1 = d.nparam > 1 ? get_param(T, di, node, c, 1, -1, 0.0) : 0.0;
  double u, u2;
  switch (d.dist) {
    case NIMB_DNORM: normal2(key, 0, u, u2); nv = p0 + p1 * u; return true;
    case NIMB_DGAMMA: nv = rgamma1(key, p0) * p1; return true;
    case NIMB_DBETA: nv = rbeta1(key, p0, p1); return true;
    case NIMB_DEXP: unif2(key, 0, u, u2); nv = -det_log(u) / p0; return true;
    case NIMB_DLNORM: normal2(key, 0, u, u2); nv = det_exp(p0 + p1 * u); return true;
    case NIMB_DLOGIS: unif2(key, 0, u, u2); nv = p0 + p1 * (det_log(u) - det_log(1.0 - u)); return true;
    case NIMB_DBINOM: nv = rbinom_inv(key, p0, p1); return !isnan(nv);
    case NIMB_DPOIS: nv = rpois_inv(key, p0); return !isnan(nv);
  }
  return false;
}
__global__ void k_init_simulate(DevTab T, int di, uint64_t seed, int* fail_flag) {
  const nimb200_decl& d = T.decl[di];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)d.n * T.C) return;
  int c = (int)(gid % T.C), node = (int)(gid / T.C);
  double* sv = T.state + (size_t)(d.value_ref + node) * T.C + c;
  if (!isnan(*sv)) return;
  DrawKey key; key.seed = seed; key.chain = (uint32_t)(T.first_chain + c); key.iter = 0;
  key.sampler = 0x7fff0000u + (uint32_t)di; key.target = (uint32_t)node;
  double nv;
  if (simulate_node(T, di, node, c, key, nv)) *sv = nv; else atomicExch(fail_flag, di + 1);
}

// online WAIC (MCMC_WAIC.R:246-268): one thread per data node walks the chains of this iteration in order and updates the node's
// running maximum / scaled sum (log-sum-exp) and Welford mean / sum of squares.  acc: [4][n] (max, sum, mean, ssq).
__global__ void k_waic_update(DevTab T, int di, double* __restrict__ acc, long long iter0, long long nburnin, long long base_count) {
  const long long it = (long long)T.ctr[0] - iter0;
  if (it <= nburnin) return;
  const int n = T.decl[di].n;
  int node = blockIdx.x * blockDim.x + threadIdx.x;
  if (node >= n || node < T.shard_lo[di] || node >= T.shard_hi[di]) return;   // data-sharded: this rank's slice of the data nodes
  double mx = acc[node], cs = acc[(size_t)n + node], mean = acc[2 * (size_t)n + node], ss = acc[3 * (size_t)n + node];
  long long cnt = base_count + (it - nburnin - 1) * (long long)T.C;
  for (int c = 0; c < T.C; c++) {
    const double lp = calc_node(T, di, node, c, -1, 0.0);
    cnt++;
    if (cnt == 1) { mx = lp; cs = 1.0; }
    else if (lp > mx) { cs = cs * exp(-lp + mx) + 1.0; mx = lp; }
    else cs = cs + exp(lp - mx);
    const double d1 = lp - mean;
    mean = mean + d1 / (double)cnt;
    const double d2 = lp - mean;
    ss = ss + d1 * d2;
  }
  acc[node] = mx; acc[(size_t)n + node] = cs; acc[2 * (size_t)n + node] = mean; acc[3 * (size_t)n + node] = ss;
}

// ------------------------------------------------------------------------------------------------ wide kernels
// (node tile x chain tile) evaluation of instances [lo,hi) (through idx if given) of one declaration.
// MODE 0: per-chain partial sums of log-densities -> partials[tile][C]
// MODE 1: per-node log-densities -> out_nodes[(k-lo)*C + c]
// MODE 2: conjugate contributions (two sums)     -> partials[tile][2][C]
template <int MODE>
__global__ void k_eval(DevTab T, int di, int lo, int hi, const int32_t* idx, int cw_log2, int npt, double* partials,
                       double* out_nodes, nimb200_dep dp, int slot, int gam) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  extern __shared__ double sh[];
  int cw = 1 << cw_log2, nl = blockDim.x >> cw_log2;
  int lane_c = threadIdx.x & (cw - 1), lane_n = threadIdx.x >> cw_log2;
  int c = blockIdx.y * cw + lane_c;
  int tile0 = lo + blockIdx.x * (nl * npt);
  double a0 = 0, a1 = 0;
  if (c < T.C) {
    Hoist H;
    if (MODE != 2) H = make_hoist(T, di, c, -1, 0.0);
    bool hp = false;
    double par = 0.0;
    if (MODE == 2) par = conj_par(T, dp, 0, c, gam, &hp);
    for (int q = 0; q < npt; q++) {
      int k = tile0 + q * nl + lane_n;
      if (k >= hi) break;
      int node = idx ? idx[k] : k;
      if (idx && (node < T.shard_lo[di] || node >= T.shard_hi[di])) continue;
      if (MODE == 0) a0 += calc_node_h(T, di, node, c, -1, 0.0, H);
      else if (MODE == 1) out_nodes[(size_t)(k - lo) * T.C + c] = calc_node_h(T, di, node, c, -1, 0.0, H);
      else conj_contrib(T, dp, node, c, slot, gam, a0, a1, hp, par);
    }
  }
  if (MODE == 1) return;
  sh[threadIdx.x] = a0;
  if (MODE == 2) sh[blockDim.x + threadIdx.x] = a1;
  __syncthreads();
  if (lane_n == 0 && c < T.C) {
    double s0 = 0, s1 = 0;
    for (int j = 0; j < nl; j++) {
      s0 += sh[(j << cw_log2) + lane_c];
      if (MODE == 2) s1 += sh[blockDim.x + (j << cw_log2) + lane_c];
    }
    if (MODE == 0) partials[(size_t)blockIdx.x * T.C + c] = s0;
    else {
      partials[((size_t)blockIdx.x * 2) * T.C + c] = s0;
      partials[((size_t)blockIdx.x * 2 + 1) * T.C + c] = s1;
    }
  }
}

// fixed-order reduction of the tile partials: dst[k][c] = sum_b partials[b][k][c].  One warp per (chain, k): lane l adds
// tiles l, l+32, ... in order, then a butterfly over the lanes -- the same order every run, so sums are reproducible.
__global__ void k_reduce(const double* __restrict__ partials, int nb, int K, int C, double* dst0, double* dst1) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= C * K) return;
  int c = w % C, kk = w / C;
  double s = 0;
  for (int b = lane; b < nb; b += 32) s += partials[((size_t)b * K + kk) * C + c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) (kk == 0 ? dst0 : dst1)[c] = s;
}

// ---- i.i.d. gamma declarations with chain-level parameters:  x[i] ~ dgamma(shape, scale),  x latent, i = 1..n
// (the random-effect layer of configs[4]: lam[g] ~ dgamma(a, b)).  Its block sum and its conjugate contributions depend on the n
// values only through  SL = sum log x  and  SM = sum x:
//   sum_i dgamma(x_i) = n (-shape log scale - lgamma(shape)) + (shape - 1) SL - SM / scale,
//   contributions to a dgamma prior on the rate (MCMC_conjugacy.R:85): (n shape, coeff SM).
// k_gamma_stats makes one pass over the values when they have changed (fixed order: lane = chain, chunk partials, k_reduce);
// k_gamma_from_stats serves every evaluation until the next change -- the RW step on the shape (two evaluations) and the
// conjugate update of the rate cost O(1) per chain instead of three passes with one logarithm per node each.
// A chain with a non-positive / non-finite value, or parameters outside the closed form's range, walks the nodes exactly as
// k_eval does.
static const int GS_ROWS = 256;
__global__ void __launch_bounds__(256) k_gamma_stats(DevTab T, int di, double* __restrict__ partials) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  __shared__ double red[2][8][32];
  __shared__ __align__(16) MathTab tab;
  load_mathtab(&tab, threadIdx.x, 256);
  __syncthreads();
  const nimb200_decl& d = T.decl[di];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int r0 = blockIdx.y * GS_ROWS, r1 = min(d.n, r0 + GS_ROWS);
  double sl = 0.0, sm = 0.0;
  if (c < T.C) {
    const double* __restrict__ xp = T.state + (size_t)d.value_ref * T.C + c;
    int r = r0 + warp;
    for (; r + 24 < r1; r += 32) {   // four loads and four logarithms in flight; added in row order as below
      double x4[4], l4[4];
#pragma unroll
      for (int j = 0; j < 4; j++) x4[j] = xp[(size_t)(r + 8 * j) * T.C];
#pragma unroll
      for (int j = 0; j < 4; j++)   // table-driven logarithm for positive normal values (<= 1 ulp), the library's otherwise
        l4[j] = (x4[j] >= 1e-300 && x4[j] < 1e300) ? log_tab(x4[j], &tab) : log(x4[j]);
#pragma unroll
      for (int j = 0; j < 4; j++) { sl += l4[j]; sm += x4[j]; }
    }
    for (; r < r1; r += 8) {
      const double x = xp[(size_t)r * T.C];
      sl += log(x);
      sm += x;
    }
  }
  red[0][warp][lane] = sl; red[1][warp][lane] = sm;
  __syncthreads();
  if (warp < 2 && c < T.C) {
    double s = 0.0;
    for (int w = 0; w < 8; w++) s += red[warp][w][lane];
    partials[((size_t)blockIdx.y * 2 + warp) * T.C + c] = s;
  }
}
// MODE 0: block sum -> dst0.  MODE 2: conjugate contributions of the n dependents -> (dst0, dst1).  One warp per chain: the
// closed form is lane 0's; a chain outside it (a proposed shape beyond 64 happens every few iterations while the RW scale is
// unadapted) has its 32 lanes stride over the nodes, fixed-order butterfly at the end -- one thread walking 1e4 nodes through
// nmath's dgamma took 10 ms.
template <int MODE>
__global__ void k_gamma_from_stats(DevTab T, int di, const double* __restrict__ stats, nimb200_dep dp, int slot, int gam,
                                   double* __restrict__ dst0, double* __restrict__ dst1) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (c >= T.C) return;
  const nimb200_decl& d = T.decl[di];
  const double SL = stats[c], SM = stats[(size_t)T.C + c];
  const bool fin = isfinite(SL) && isfinite(SM);
  if (MODE == 0) {
    const Hoist H = make_hoist(T, di, c, -1, 0.0);
    if (H.gfast && fin) {
      if (lane == 0) dst0[c] = fma(H.p0 - 1.0, SL, (double)d.n * H.l1) - SM / H.p1;
      return;
    }
    double a = 0.0;
    for (int i = lane; i < d.n; i += 32) a += calc_node_h(T, di, i, c, -1, 0.0, H);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (lane == 0) dst0[c] = a;
  } else {
    bool hp = false;
    const double par = conj_par(T, dp, 0, c, gam, &hp);
    double coeff = 1.0;
    if (dp.link == NIMB_LINK_MULTIPLICATIVE) coeff = get_param(T, dp.decl, 0, c, dp.param, slot, 1.0);
    if (fin && hp) {
      if (lane == 0) { dst0[c] = (double)d.n * par; dst1[c] = coeff * SM; }
      return;
    }
    double a0 = 0.0, a1 = 0.0;
    for (int i = lane; i < d.n; i += 32) conj_contrib(T, dp, i, c, slot, gam, a0, a1, hp, par);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o); }
    if (lane == 0) { dst0[c] = a0; dst1[c] = a1; }
  }
}

// wide RW: proposal for every chain; the old value is parked in tsave
// prior_new (may be null): the target's own log-density under the proposal, i.e. what k_eval<0> would return for the
// sampler's first dependency entry (one node) -- the same hoisted evaluation, one launch less per scalar RW step
__global__ void k_rw_propose(DevTab T, int sid, uint64_t seed, uint64_t iter, SamplerDev S, double* tsave, double* pls,
                             double* prior_new, int prior_node) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= T.C) return;
  const nimb200_sampler sp = T.samplers[sid];
  DrawKey key = make_key(T, seed, iter, sid, 0, c);
  double* sv = T.state + (size_t)sp.target_slot * T.C + c;
  double z, z1, propLogScale;
  normal2(key, 0, z, z1);
  double cur = *sv;
  tsave[c] = cur;
  *sv = rw_proposal(sp, cur, S.scale[c], z, propLogScale);
  pls[c] = propLogScale;
  if (prior_new) {
    const nimb200_dep d0 = T.deps[sp.dep_off];
    const Hoist H = make_hoist(T, d0.decl, c, -1, 0.0);
    prior_new[c] = 0.0 + calc_node_h(T, d0.decl, prior_node, c, -1, 0.0, H);
  }
}

// wide RW_block: rmnorm_chol(values, chol_propCov_scale, prec_param = 0), dists.cpp:1132-1164.
// Block = 16 chains x 64 rows: thread (c, j) first draws the normals of pair j, then forms row j of x := U^T z with the
// same descending accumulation order as the reference-BLAS dtrmv('U','T','N') loop, so the proposal is bit-identical to
// the sequential version while 16 x d threads share the work and the U loads stay coalesced over chains.
__global__ void k_block_propose(DevTab T, int sid, uint64_t seed, uint64_t iter, int tr, SamplerDev S, double* tsave) {
  __shared__ double z[NIMB200_MAX_BLOCK_D][16];
  const int cl = threadIdx.x, j = threadIdx.y;
  const int c = blockIdx.x * 16 + cl;
  const nimb200_sampler sp = T.samplers[sid];
  const int d = sp.d, npair = (d + 1) / 2;
  if (c < T.C && j < npair) {
    DrawKey key = make_key(T, seed, iter, sid, 0, c);
    double z0, z1;
    normal2(key, tr * (npair + 1) + j, z0, z1);
    z[2 * j][cl] = z0;
    if (2 * j + 1 < d) z[2 * j + 1][cl] = z1;
  }
  __syncthreads();
  if (c >= T.C || j >= d) return;
  const double* U = S.cholScale;
  double temp = z[j][cl] * U[(size_t)(j + j * d) * T.C + c];
  for (int i = j - 1; i >= 0; i--) temp += U[(size_t)(i + j * d) * T.C + c] * z[i][cl];
  double* sv = T.state + (size_t)(sp.target_slot + j) * T.C + c;
  double cur = *sv;
  tsave[(size_t)j * T.C + c] = cur;
  *sv = temp + cur;
}

// wide MH accept/reject + commit.  newsum[e][C] / oldsum = depcache rows of this sampler's dependency entries.
// Block = 32 chains (x) by MHF_J lanes (y): lane 0 of a chain takes the decision, then the chain's lanes share the per-element
// work of a block sampler -- putting the old values back after a rejection and recording the iteration for the covariance
// adaptation (one thread per chain walked 2 x 50 dependent loads and stores: 29 us per step at any chain count).
static const int MHF_J = 16;
__global__ void __launch_bounds__(32 * MHF_J)
k_mh_finish(DevTab T, int sid, uint64_t seed, uint64_t iter, int call, SamplerDev S, const double* newsum,
            double* depcache, const int32_t* __restrict__ dep_slot, const double* tsave, const double* pls,
            int* __restrict__ accept_out) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  __shared__ int s_jump[32], s_ran[32];
  const int cx = threadIdx.x, jy = threadIdx.y;
  const int c = blockIdx.x * 32 + cx;
  const bool live = c < T.C;
  const nimb200_sampler sp = T.samplers[sid];
  const int d = sp.type == NIMB_SAMPLER_RW_BLOCK ? sp.d : 1;
  if (jy == 0 && live) {
    DrawKey key = make_key(T, seed, iter, sid, 0, c);
    double lp0 = 0, lp1 = 0;
    for (int e = 0; e < sp.n_dep; e++) {
      double diff = newsum[(size_t)e * T.C + c] - depcache[(size_t)dep_slot[sp.dep_off + e] * T.C + c];
      if (T.deps[sp.dep_off + e].stage == 0) lp0 += diff; else lp1 += diff;
    }
    bool jump;
    if (lp0 == -INFINITY) jump = false;
    else jump = decide(lp0 + lp1 + (pls ? pls[c] : 0.0), key, call);
    if (accept_out) accept_out[c] = jump ? 1 : 0;
    if (jump)
      for (int e = 0; e < sp.n_dep; e++) depcache[(size_t)dep_slot[sp.dep_off + e] * T.C + c] = newsum[(size_t)e * T.C + c];
    int ran = 0;
    if (sp.adaptive) {
      if (sp.type == NIMB_SAMPLER_RW) rw_adapt(sp, S, c, jump, T.C);
      else {  // RW_block adaptiveProcedure, first half (MCMC_samplers.R:716-722); the covariance step is k_block_adapt
        ran = S.timesRan[c] + 1;
        S.timesRan[c] = ran;
        if (jump) S.timesAccepted[c] += 1;
      }
    }
    s_jump[cx] = jump ? 1 : 0;
    s_ran[cx] = ran;
  }
  __syncthreads();
  if (!live) return;
  const bool jump = s_jump[cx] != 0;
  const bool record = sp.adaptive && sp.type == NIMB_SAMPLER_RW_BLOCK && !sp.adapt_scale_only;
  const int ran = s_ran[cx];
  for (int j = jy; j < d; j += MHF_J) {
    double* sv = T.state + (size_t)(sp.target_slot + j) * T.C + c;
    double v;
    if (jump) v = *sv;                                  // the proposal stays
    else { v = tsave[(size_t)j * T.C + c]; *sv = v; }   // the old value comes back
    if (record) S.empirSamp[((size_t)(ran - 1) + (size_t)j * sp.adapt_interval) * T.C + c] = v;
  }
}

// RW_block adaptation every adaptInterval iterations: one block per chain (MCMC_samplers.R:723-756,
// calcAdaptationFactor MCMC_utils.R:177-205).  Loop orders match the oracle so results are bit-identical.
__global__ void k_block_adapt(DevTab T, int sid, SamplerDev S) {
  __shared__ double U[NIMB200_MAX_BLOCK_D * NIMB200_MAX_BLOCK_D];  // propCov, overwritten in place by its upper factor
  __shared__ double s_gamma1, s_scale, s_diag;
  __shared__ int s_adapted;
  const nimb200_sampler sp = T.samplers[sid];
  int c = blockIdx.x, d = sp.d, AI = sp.adapt_interval, C = T.C;
  int tid = threadIdx.x, nt = blockDim.x;
  if (tid == 0) {
    const double optAR[5] = {0.44, 0.35, 0.32, 0.25, 0.234};
    int ran = S.timesRan[c];
    double acceptanceRate = (double)S.timesAccepted[c] / (double)ran;
    int adapted = S.timesAdapted[c] + 1;
    S.timesAdapted[c] = adapted;
    s_adapted = adapted;
    if (S.hist && adapted <= S.hist_cap) {  // MCMC_samplers.R:726-739
      S.hist[((size_t)(adapted - 1) * S.hist_K + 0) * C + c] = S.scale[c];
      S.hist[((size_t)(adapted - 1) * S.hist_K + 1) * C + c] = acceptanceRate;
    }
    double gamma1 = 1 / pow((double)adapted + 3, sp.adapt_factor_exponent);
    double adaptFactor = exp(10 * gamma1 * (acceptanceRate - optAR[d > 5 ? 4 : d - 1]));
    s_gamma1 = gamma1;
    s_scale = S.scale[c] * adaptFactor;
    S.scale[c] = s_scale;
  }
  __syncthreads();
  if (S.hist && S.hist_K > 2 && s_adapted <= S.hist_cap)   // propCovHistory: the covariance this interval's proposals were drawn with
    for (int e = tid; e < d * d; e += nt) S.hist[((size_t)(s_adapted - 1) * S.hist_K + 2 + e) * C + c] = S.propCov[(size_t)e * C + c];
  if (!sp.adapt_scale_only) {
    for (int j = tid; j < d; j += nt) {  // centre the columns of empirSamp
      double mu = 0;
      for (int r = 0; r < AI; r++) mu += S.empirSamp[((size_t)r + (size_t)j * AI) * C + c];
      mu = mu / AI;
      for (int r = 0; r < AI; r++) S.empirSamp[((size_t)r + (size_t)j * AI) * C + c] -= mu;
    }
    __syncthreads();
    int ran = S.timesRan[c];
    for (int e = tid; e < d * d; e += nt) {  // propCov += gamma1 * (E'E/(n-1) - propCov)
      int a = e % d, b = e / d;
      double acc = 0;
      for (int r = 0; r < AI; r++)
        acc += S.empirSamp[((size_t)r + (size_t)a * AI) * C + c] * S.empirSamp[((size_t)r + (size_t)b * AI) * C + c];
      double ec = acc / (double)(ran - 1);
      double pc = S.propCov[(size_t)e * C + c];
      pc = pc + s_gamma1 * (ec - pc);
      S.propCov[(size_t)e * C + c] = pc;
      U[e] = pc;
    }
    __syncthreads();
    for (int j = 0; j < d; j++) {  // column Cholesky, upper factor
      if (tid == 0) {
        double s = U[j + j * d];
        for (int k = 0; k < j; k++) s -= U[k + j * d] * U[k + j * d];
        s = sqrt(s);
        U[j + j * d] = s;
        s_diag = s;
      }
      __syncthreads();
      for (int i = j + 1 + tid; i < d; i += nt) {
        double v = U[j + i * d];
        for (int k = 0; k < j; k++) v -= U[k + j * d] * U[k + i * d];
        U[j + i * d] = v / s_diag;
      }
      __syncthreads();
    }
    for (int e = tid; e < d * d; e += nt) S.chol[(size_t)e * C + c] = (e % d) <= (e / d) ? U[e] : 0.0;
    __syncthreads();
  }
  for (int e = tid; e < d * d; e += nt) S.cholScale[(size_t)e * C + c] = S.chol[(size_t)e * C + c] * s_scale;
  if (tid == 0) { S.timesRan[c] = 0; S.timesAccepted[c] = 0; }
}

// wide conjugate: draw from the posterior given the reduced contributions
__global__ void k_conj_finish(DevTab T, int sid, uint64_t seed, uint64_t iter, const double* contrib /*[n_dep][2][C]*/) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= T.C) return;
  const nimb200_sampler sp = T.samplers[sid];
  const int gam = conj_kind(sp.type);
  DrawKey key = make_key(T, seed, iter, sid, 0, c);
  const nimb200_dep tp = T.deps[sp.dep_off];
  int tnode = tp.a;  // stage-0 entry of a single-target sampler: RANGE(node, node+1)
  double prior0 = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
  double prior1 = get_param(T, tp.decl, tnode, c, gam == 2 ? 1 : (gam == 1 ? NIMB_PARAM_RATE : NIMB_PARAM_TAU), -1, 0.0);
  double c0 = 0, c1 = 0;
  for (int e = 0; e < sp.n_dep; e++) {
    if (T.deps[sp.dep_off + e].stage != 1) continue;
    c0 += contrib[((size_t)e * 2) * T.C + c];
    c1 += contrib[((size_t)e * 2 + 1) * T.C + c];
  }
  double nv;
  if (gam == 2) {
    nv = rbeta1(key, prior0 + c0, prior1 + c1);
  } else if (gam == 1) {
    double shape = prior0 + c0, scale = 1 / (prior1 + c1);
    nv = rgamma1(key, shape) * scale;
  } else {
    double mean = (prior0 * prior1 + c0) / (prior1 + c1), sd = pow(prior1 + c1, -0.5), z, z1;
    normal2(key, 0, z, z1);
    nv = mean + sd * z;
  }
  T.state[(size_t)sp.target_slot * T.C + c] = nv;
}

__global__ void k_sampler_reset(DevTab T, int sid, SamplerDev S, const double* propcov0 /*d*d col-major or null*/) {
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int TT = sp.type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp.n_targets;
  if (gid >= (size_t)TT * T.C) return;
  S.scale[gid] = sp.scale;
  S.timesRan[gid] = 0; S.timesAccepted[gid] = 0; S.timesAdapted[gid] = 0;
  if (sp.type == NIMB_SAMPLER_SLICE) S.propCov[gid] = 0.0;  // sumJumps
}

// chol(propCov) at reset: one block per chain (propCov is the same for all chains at reset)
__global__ void k_block_reset(DevTab T, int sid, SamplerDev S, const double* propcov0) {
  __shared__ double U[NIMB200_MAX_BLOCK_D * NIMB200_MAX_BLOCK_D];
  __shared__ double s_diag;
  const nimb200_sampler sp = T.samplers[sid];
  int c = blockIdx.x, d = sp.d, C = T.C, tid = threadIdx.x, nt = blockDim.x;
  for (int e = tid; e < d * d; e += nt) {
    U[e] = propcov0 ? propcov0[e] : ((e % d) == (e / d) ? 1.0 : 0.0);
    S.propCov[(size_t)e * C + c] = U[e];
  }
  __syncthreads();
  for (int j = 0; j < d; j++) {
    if (tid == 0) {
      double s = U[j + j * d];
      for (int k = 0; k < j; k++) s -= U[k + j * d] * U[k + j * d];
      s = sqrt(s);
      U[j + j * d] = s;
      s_diag = s;
    }
    __syncthreads();
    for (int i = j + 1 + tid; i < d; i += nt) {
      double v = U[j + i * d];
      for (int k = 0; k < j; k++) v -= U[k + j * d] * U[k + i * d];
      U[j + i * d] = v / s_diag;
    }
    __syncthreads();
  }
  for (int e = tid; e < d * d; e += nt) {
    double u = (e % d) <= (e / d) ? U[e] : 0.0;
    S.chol[(size_t)e * C + c] = u;
    S.cholScale[(size_t)e * C + c] = u * sp.scale;
  }
}

// mvSamples row: nimCopy(model -> mvSamples, nodes = monitors), MCMC_build.R:294-303
__global__ void k_monitor(DevTab T, const int32_t* mon, int nmon, double* row /*[nmon][C]*/) {
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)nmon * T.C) return;
  int c = (int)(gid % T.C), j = (int)(gid / T.C);
  row[gid] = T.state[(size_t)mon[j] * T.C + c];
}

// graph-replayable form: decides from the device iteration counter whether this iteration is saved and where
// (iteration it = ctr[0] - iter0 of this run; rows_before rows were already in the buffer)
// The last block to finish also advances the counter (every block has read it by then), which saves the separate launch.
// what changes from one nimb200_mcmc_run call to the next lives in device memory, not in kernel arguments, so that the captured
// iteration graph of one call can be replayed by the next (a `reset = FALSE` continuation does not pay for a new capture)
struct RunParams { long long rows_before, iter0, nburnin; int thin, pad; double* samples; };

__global__ void k_monitor_ctr(DevTab T, const int32_t* mon, int nmon, const RunParams* __restrict__ rp, unsigned long long* ctr,
                              unsigned int* ticket) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  const long long rows_before = rp->rows_before, iter0 = rp->iter0, nburnin = rp->nburnin;
  const int thin = rp->thin;
  double* samples = rp->samples;
  long long it = (long long)T.ctr[0] - iter0;
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (it > nburnin && (it - nburnin) % thin == 0 && gid < (size_t)nmon * T.C) {
    int c = (int)(gid % T.C), j = (int)(gid / T.C);
    long long row = rows_before + (it - nburnin) / thin - 1;
    samples[(size_t)row * nmon * T.C + gid] = T.state[(size_t)mon[j] * T.C + c];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(ticket, 1u) == gridDim.x - 1) { *ticket = 0; ctr[0] += 1; }
  }
}
__global__ void k_bump_iter(unsigned long long* ctr) { ctr[0] += 1; }

// [rows][nmon][C] -> [C][rows][nmon]
__global__ void k_samples_out(const double* src, double* dst, int64_t rows, int nmon, int C) {
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)rows * nmon * C;
  if (gid >= total) return;
  int c = (int)(gid % C);
  size_t rj = gid / C;
  dst[(size_t)c * rows * nmon + rj] = src[gid];
}

// [C][n] host layout <-> [n][C] device layout
__global__ void k_transpose(const double* src, double* dst, int rows, int cols) {  // src[rows][cols] -> dst[cols][rows]
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)rows * cols) return;
  int r = (int)(gid / cols), q = (int)(gid % cols);
  dst[(size_t)q * rows + r] = src[gid];
}

__global__ void k_draws(uint64_t seed, int chain, uint64_t iter, int sampler, int target, int kind, double a, int n, double* out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  DrawKey key; key.seed = seed; key.chain = (uint32_t)chain; key.iter = iter; key.sampler = (uint32_t)sampler; key.target = (uint32_t)target;
  double x0, x1;
  if (kind == 0) { normal2(key, k / 2, x0, x1); out[k] = (k & 1) ? x1 : x0; }
  else if (kind == 1) { unif2(key, k / 2, x0, x1); out[k] = (k & 1) ? x1 : x0; }
  else if (kind == 2) { key.target = (uint32_t)(target + k); out[k] = rgamma1(key, a); }
  else if (kind == 3) { key.target = (uint32_t)(target + k); out[k] = rbeta1(key, a, 1.0 + 0.5 * a); }
  else if (kind == 4) { key.target = (uint32_t)(target + k); out[k] = rpois_inv(key, a); }
  else { key.target = (uint32_t)(target + k); out[k] = rbinom_inv(key, 25.0, a); }
}

// ------------------------------------------------------------------------------------------------ host model
struct nimb200_model {
  int device = 0, C = 0, first_chain = 0;
  nimb200_table t{};  // host copy (arrays below)
  std::vector<nimb200_op> ops;
  std::vector<nimb200_decl> decl;
  std::vector<nimb200_sampler> samplers;
  std::vector<nimb200_dep> deps;
  std::vector<int32_t> monitors;
  std::vector<int64_t> darr_len, iarr_len;
  std::vector<double*> d_darr;
  std::vector<int32_t*> d_iarr;
  std::vector<std::vector<int32_t>> h_iarr;  // host copies of the int arrays (segment offsets are read when launches are sized)
  DevTab T{};
  void* d_ops = nullptr; void* d_decl = nullptr; void* d_samplers = nullptr; void* d_deps = nullptr;
  void* d_darr_tab = nullptr; void* d_iarr_tab = nullptr;
  int32_t* d_mon = nullptr;
  double* d_state = nullptr;
  double* d_stage = nullptr;    // staging for host<->device transposes, n_slots*C
  double* d_lpcache = nullptr;  // [n_decl][C] block sums behind calculate/calculateDiff/getLogProb
  double* d_depcache = nullptr; // [n_deps][C] block sums of the CURRENT state per dependency entry
  double* d_newsum = nullptr;   // [max_dep][2][C]
  double* d_tsave = nullptr;    // [MAX_BLOCK_D][C]
  double* d_pls = nullptr;      // [C]
  double* d_partials = nullptr; // scratch: tile partial sums
  size_t partials_cap = 0;
  std::vector<SamplerDev> sdev;
  std::vector<char> wide;       // per sampler: wide path?
  std::vector<char> dep_valid;  // per cache slot: depcache row valid?
  std::vector<char> lp_valid;   // per declaration: d_lpcache row holds the block sum of the CURRENT state (set by calculate() on the
                                // whole declaration, cleared by anything that changes the chain state)
  std::vector<int32_t> dep_slot; // dependency entry -> cache slot: entries of wide samplers with the same (declaration, selection) share one
  int32_t* d_dep_slot = nullptr;
  std::vector<std::vector<int>> invalidates;  // per sampler: dependency entries (of any sampler) it invalidates
  std::vector<int> ran;         // RW_block: iterations since last adaptation
  std::vector<double*> d_xpad;  // per decl: padded X for the fused family (or null)
  std::vector<void*> d_nodeconst;    // per decl: packed nodes (PoisNode: y, x, data-only constant, lam index) of the fused Poisson family, or null
  std::vector<char> pois_sweep;      // per sampler: conjugate sweep served by k_pois_conj_sweep
  std::vector<char> bern_ok;         // per decl: Bernoulli-logit-linear family usable
  std::vector<char> bern_sweep;      // per sampler: binary sweep served by k_bernlin_binary_sweep
  std::vector<std::array<int, 2>> bern_byp_slot;  // per such sweep: cache slots of the whole-declaration block sums (own declaration,
  std::vector<double*> d_bern_contrib;            // dependents) it can refresh, or -1; [2][n_targets][C] per-(site, chain) shares
  std::vector<char> norm_ok;         // per decl: streaming normal family usable
  std::vector<int> norm_vspan;       // per decl: number of elements of the gathered vector that are reachable
  std::vector<char> norm_sweep;      // per sampler: conjugate sweep served by k_norm_group_sums + k_norm_conj_draw
  std::vector<double*> d_pois_ysum;  // per Poisson conjugate sweep: [4][n_targets] data-only statistics of each target's nodes (sums of y, A, x y; max |x|)
  std::vector<int> pois_byp_slot;    // per such sweep: cache slot of the dependents' whole-declaration block sum it can refresh, or -1
  std::vector<double*> d_pois_contrib; // [n_targets][C] per-(target, chain) shares of that block sum
  std::vector<char> pois_ysum_valid;
  // group sums S[t][c] = sum over target t's nodes of exp(b1 x) for chain c's CURRENT b1 (fused_pois.cuh): with them a block sum or
  // a conjugate sweep of the Poisson family reads no node; one pass (k_pois_conj_sweep<3>) when b1 moves
  std::vector<double*> d_pois_S, d_pois_Snew;
  std::vector<int32_t*> d_pois_list;   // per sweep: chains that need the node pass [C] + their count
  std::vector<double*> d_pois_chainpar;   // per sweep: [3][C] exp(b0), prior shape, prior rate (k_pois_chain_params)
  std::vector<char> pois_chain_level;     // ... usable: the targets' prior parameter programs are chain-level
  std::vector<int32_t> h_pconst;          // host copy of T.pconst
  std::vector<char> pois_S_valid, pois_S_pending;
  // i.i.d. gamma declarations with chain-level parameters (k_gamma_stats): eligibility, [2][C] statistics, validity
  std::vector<char> gam_iid, gam_stats_valid;
  std::vector<double*> d_gam_stats;
  bool pois_series = true;   // group sums from the data-only moment series (fused_pois.cuh); NIMB200_NO_SERIES=1 keeps the node pass
  SmallPlan small{};                // whole-iteration kernel for small all-sweep models (k_small_iters); small.n == 0: not eligible
  SamplerDev* d_sdev = nullptr;     // device copy of sdev[] for it
  int* d_accept = nullptr;        // [C] accept flags of the last wide MH decision
  int prop_lo = -1, prop_hi = -1; // state slots that currently hold a proposal (between propose and finish)
  std::vector<double*> d_norm_gs;    // per such sampler: [n_targets][2] (sum y, count) of every target's nodes -- data only, so it
  std::vector<char> norm_gs_valid;   // is computed (and, data-sharded, all-reduced) once and kept until the shards change
  std::vector<double*> d_norm_gss;   // per such sampler: [n_targets] centred sum of squares of every target's nodes (data only too)
  std::vector<char> norm_gss_ok;     // ... all finite and every node of the dependent declaration covered: block sums may use it
  double* d_normtmp = nullptr;       // [2][C]
  // data-node sharding (nimb200_enable_data_sharding)
  int rank = 0, nranks = 1;
  std::vector<int32_t> shard_lo, shard_hi, sharded;
  int32_t *d_shard_lo = nullptr, *d_shard_hi = nullptr, *d_sharded = nullptr;
  void* nccl_comm = nullptr;
  // peer exchange (nimb200_peer_export / nimb200_enable_data_sharding_peer): device-side all-reduce over NVLink peer memory
  PeerXchg xchg{};
  bool xchg_on = false;
  void* xchg_local = nullptr;               // this rank's exchange allocation
  void* xchg_mapped[XCHG_MAX_RANKS] = {nullptr};  // peers' allocations opened through CUDA IPC (to be closed)
  int64_t allreduces = 0;                   // all-reduces enqueued by the last run
  double* d_sweepbuf = nullptr;  // [T][3][C] partial sums of a sharded sweep
  size_t sweepbuf_cap = 0;
  std::vector<char> sweep_sharded;  // per sampler: touches a sharded declaration
  std::vector<char> target_sharded; // per sampler: conjugate sweep whose targets are OWNED by ranks (shards cut on group boundaries)
  std::vector<size_t> sweep_deps;   // per sampler: dependent nodes of target 0 (execution-shape heuristic)
  unsigned long long* d_ctr = nullptr;  // device iteration counter (DevTab::ctr)
  unsigned int* d_ticket = nullptr;     // block ticket of k_monitor_ctr (the last block advances the counter)
  cudaGraphExec_t graph_exec = nullptr;
  // the captured iteration stays usable across runs as long as what was baked into it still holds
  std::vector<char> graph_pattern;      // cache-validity pattern at the start of the captured iteration
  uint64_t graph_seed = 0;
  int64_t graph_launches = 0, graph_allreduces = 0;
  void* d_runp = nullptr;               // RunParams
  double* d_samples = nullptr;  // [cap][nmon][C]
  int64_t rows = 0, cap = 0, iter_done = 0;
  bool waic = false;                 // online WAIC enabled
  std::vector<double*> d_waic;       // per data declaration: [4][n] accumulators (null for latent declarations)
  int64_t waic_count = 0;            // (iteration, chain) samples accumulated so far
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool time_samplers = false;            // run(time = TRUE): per-sampler CUDA-event timing
  std::vector<cudaEvent_t> ev_s;         // 2 events per sampler
  std::vector<double> sampler_ms;        // accumulated per sampler
  double last_ms = 0;
  int64_t last_launches = 0, launches = 0;
  std::vector<void*> allocs;
};

// ------------------------------------------------------------------------------------------------ NCCL (dlopen)
// Only the data-sharded mode needs a collective: one all-reduce (sum, FP64) of per-chain partial sums per block-sum
// evaluation.  NCCL is loaded at run time (the copy torch already mapped if there is one) so that the library has no
// link-time dependency on it and single-GPU users never touch it.
#include <dlfcn.h>
#include <unistd.h>
namespace {
struct NcclId { char internal[128]; };
typedef int (*fn_getid)(NcclId*);
typedef int (*fn_init)(void**, int, NcclId, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);
struct NcclApi {
  void* h = nullptr;
  fn_getid getid = nullptr; fn_init init = nullptr; fn_allreduce allreduce = nullptr; fn_destroy destroy = nullptr;
  fn_errstr errstr = nullptr;
};
NcclApi g_nccl;
int load_nccl() {
  if (g_nccl.h) return 0;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    g_nccl.h = dlopen(nm, RTLD_NOW | RTLD_NOLOAD);
    if (g_nccl.h) break;
  }
  for (int i = 0; !g_nccl.h && i < 2; i++) g_nccl.h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  if (!g_nccl.h) return 1;
  g_nccl.getid = (fn_getid)dlsym(g_nccl.h, "ncclGetUniqueId");
  g_nccl.init = (fn_init)dlsym(g_nccl.h, "ncclCommInitRank");
  g_nccl.allreduce = (fn_allreduce)dlsym(g_nccl.h, "ncclAllReduce");
  g_nccl.destroy = (fn_destroy)dlsym(g_nccl.h, "ncclCommDestroy");
  g_nccl.errstr = (fn_errstr)dlsym(g_nccl.h, "ncclGetErrorString");
  if (!g_nccl.getid || !g_nccl.init || !g_nccl.allreduce || !g_nccl.destroy) { g_nccl.h = nullptr; return 1; }
  return 0;
}
}  // namespace

static int allreduce_sum(nimb200_model* m, double* buf, size_t count);
static inline bool comm_on(const nimb200_model* m) { return m->nccl_comm != nullptr || m->xchg_on; }

// Programmatic dependent launch: the kernel may be scheduled while its predecessor on the stream still runs; it must call
// cudaGridDependencySynchronize() before touching anything the predecessor writes (every kernel launched through here does so
// first thing, after letting its own successor be scheduled with cudaTriggerProgrammaticLaunchCompletion()).
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  static const bool off = getenv("NIMB200_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = off ? 0 : 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

static const int EVAL_THREADS = 256;
static const int EVAL_NPT = 32;

// Host -> device copy of a large pageable array through a ring of pinned staging buffers filled by several host threads: a plain
// cudaMemcpy from pageable memory runs at ~10 GB/s (one thread copying into the driver's staging buffer); with four threads
// filling 16 MiB pinned chunks while the previous chunk is in flight the copy approaches the PCIe rate.  Process-wide buffers,
// allocated on first use (64 MiB of pinned host memory), guarded by a mutex.
#include <mutex>
#include <thread>
namespace {
constexpr size_t STAGE_BYTES = 16u << 20;
constexpr int STAGE_BUFS = 4, STAGE_THREADS = 4;
struct Staging {
  std::mutex mu;
  char* buf[STAGE_BUFS] = {nullptr};
  cudaEvent_t done[STAGE_BUFS];
  bool ready = false, failed = false;
} g_stage;
}  // namespace
static cudaError_t upload_large(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  if (bytes < (32u << 20) || getenv("NIMB200_NO_STAGED_UPLOAD")) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
  std::lock_guard<std::mutex> lk(g_stage.mu);
  if (!g_stage.ready && !g_stage.failed) {
    for (int b = 0; b < STAGE_BUFS && !g_stage.failed; b++) {
      if (cudaHostAlloc((void**)&g_stage.buf[b], STAGE_BYTES, cudaHostAllocPortable) != cudaSuccess) g_stage.failed = true;
      else if (cudaEventCreateWithFlags(&g_stage.done[b], cudaEventDisableTiming) != cudaSuccess) g_stage.failed = true;
    }
    cudaGetLastError();
    g_stage.ready = !g_stage.failed;
  }
  if (!g_stage.ready) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
  size_t off = 0;
  for (int k = 0; off < bytes; k++) {
    const int b = k % STAGE_BUFS;
    const size_t n = std::min(STAGE_BYTES, bytes - off);
    if (k >= STAGE_BUFS) { cudaError_t e = cudaEventSynchronize(g_stage.done[b]); if (e != cudaSuccess) return e; }
    std::thread th[STAGE_THREADS];
    const size_t per = (n + STAGE_THREADS - 1) / STAGE_THREADS;
    for (int t = 0; t < STAGE_THREADS; t++) {
      const size_t lo = std::min(n, per * t), hi = std::min(n, per * (t + 1));
      th[t] = std::thread([=]() { if (hi > lo) memcpy(g_stage.buf[b] + lo, (const char*)src + off + lo, hi - lo); });
    }
    for (int t = 0; t < STAGE_THREADS; t++) th[t].join();
    cudaError_t e = cudaMemcpyAsync((char*)dst + off, g_stage.buf[b], n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaEventRecord(g_stage.done[b], st);
    if (e != cudaSuccess) return e;
    off += n;
  }
  return cudaStreamSynchronize(st);   // the staging buffers are free again when this returns
}

template <class Tp>
static int dev_alloc(nimb200_model* m, Tp** p, size_t count) {
  void* q = nullptr;
  CK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(Tp)));
  m->allocs.push_back(q);
  *p = (Tp*)q;
  return 0;
}

static int cw_log2_for(int C) {
  int l = 0;
  while ((1 << l) < C && l < 5) l++;
  return l;
}

// nodes per thread of k_eval: 32 for large declarations; halved until the grid has two CTAs per SM, so that a small
// declaration (a prior block of 1e4 random effects, 8 chains) is one interpreter pass per thread and not 32 in sequence
static int eval_npt(int C, int count) {
  int cwl = cw_log2_for(C), nl = EVAL_THREADS >> cwl, ct = (C + (1 << cwl) - 1) >> cwl;
  int npt = EVAL_NPT;
  while (npt > 1 && (long long)((count + nl * npt - 1) / (nl * npt)) * ct < 296) npt >>= 1;
  return npt;
}
static int eval_tiles(int C, int count) {
  int nl = EVAL_THREADS >> cw_log2_for(C);
  int per = nl * eval_npt(C, count);
  return (count + per - 1) / per;
}

static PoisArgs pois_args(nimb200_model* m, int di) {
  const nimb200_decl& d = m->decl[di];
  PoisArgs a;
  a.y = m->d_darr[d.value_ref]; a.x = m->d_darr[d.fam_arg[0]];
  a.mode = d.fam_arg[2]; a.arg = d.fam_arg[3];
  a.gidx = a.mode == NIMB_IDX_IARR ? m->d_iarr[a.arg] : nullptr;
  a.lam_slot = d.fam_arg[1]; a.b0_slot = d.fam_arg[4]; a.b1_slot = d.fam_arg[5]; a.n = d.n;
  return a;
}

static BernArgs bern_args(nimb200_model* m, int di) {
  const nimb200_decl& d = m->decl[di];
  BernArgs a;
  a.x = m->d_darr[d.fam_arg[0]];
  a.vdata = d.value_kind == 0 ? m->d_darr[d.value_ref] : nullptr;
  a.v_slot = d.value_kind == 1 ? d.value_ref : -1;
  a.c0_slot = d.fam_arg[1]; a.c1_slot = d.fam_arg[2]; a.gate_slot = d.fam_arg[3]; a.gmode = d.fam_arg[4]; a.garg = d.fam_arg[5];
  a.gidx = (a.gate_slot >= 0 && a.gmode == NIMB_IDX_IARR) ? m->d_iarr[a.garg] : nullptr;
  a.n = d.n;
  return a;
}
// node chunks for the lane = chain block-sum kernels: enough CTAs to fill the machine several times over, at most 512
static int family_chunks(int C, int count, int tile) {
  int ctiles = (C + 31) / 32;
  int want = (148 * 24 + ctiles - 1) / ctiles;
  int maxc = std::max(1, (count + tile - 1) / tile);
  return std::max(1, std::min(512, std::min(want, maxc)));
}
static NormArgs norm_args(nimb200_model* m, int di) {
  const nimb200_decl& d = m->decl[di];
  NormArgs a;
  a.y = m->d_darr[d.value_ref];
  a.a_off = d.fam_arg[0]; a.a_len = d.fam_arg[1]; a.v_slot = d.fam_arg[2]; a.mode = d.fam_arg[3]; a.arg = d.fam_arg[4];
  a.gidx = a.mode == NIMB_IDX_IARR ? m->d_iarr[a.arg] : nullptr;
  a.decl = di;
  a.a_slot = (a.a_len == 1 && m->ops[a.a_off].op == NIMB_OP_CVAR) ? m->ops[a.a_off].a : -1;
  return a;
}
static int norm_blocks(int count) {
  if (count <= 0) return 1;
  int npb = std::max(4 * NF_THREADS, ((count + 591) / 592 + 4 * NF_THREADS - 1) / (4 * NF_THREADS) * (4 * NF_THREADS));
  return (count + npb - 1) / npb;
}

// Poisson family, group-sum path: is it usable for sampler s right now (not with ungrouped data shards)?
static bool pois_S_usable(const nimb200_model* m, size_t s) {
  return m->pois_sweep[s] && m->d_pois_S[s] && !getenv("NIMB200_NO_BYPRODUCT") &&
         !(m->nranks > 1 && m->sweep_sharded[s] && !m->target_sharded[s]);
}
static int ensure_pois_ysum(nimb200_model* m, int s) {
  if (m->pois_ysum_valid[s]) return 0;
  const nimb200_sampler& sp = m->samplers[s];
  const int dd = m->deps[sp.dep_off + 1].decl;
  PoisArgs pa = pois_args(m, dd);
  CK(cudaMemsetAsync(m->d_pois_ysum[s] + pois_stats_doubles((size_t)sp.n_targets) - 1, 0, sizeof(double), m->stream));   // max over targets
  k_pois_group_ysum<<<(sp.n_targets + 127) / 128, 128, 0, m->stream>>>(m->T, s, pa, (const PoisNode*)m->d_nodeconst[dd], m->d_pois_ysum[s]);
  m->launches++;
  const bool sh = m->nranks > 1 && m->sweep_sharded[s] && !m->target_sharded[s];
  if (sh && comm_on(m) && allreduce_sum(m, m->d_pois_ysum[s], (size_t)sp.n_targets * 3)) return 1;
  m->pois_ysum_valid[s] = 1;
  return 0;
}
// one pass over the nodes: S (or Snew) [t][c] = sum_i exp(b1 x_i) under the b1 values the state holds NOW
static int launch_pois_S(nimb200_model* m, int s, double* out) {
  const nimb200_sampler& sp = m->samplers[s];
  PoisArgs pa = pois_args(m, m->deps[sp.dep_off + 1].decl);
  const dim3 grid((m->C + 31) / 32, (sp.n_targets + PS3_TPB - 1) / PS3_TPB);
  if (m->pois_series && m->d_pois_list[s]) {
    // moment series where it applies; the node pass for the listed chains only (its grid is sized for all of them: blocks
    // beyond the list return at once, so the launch sequence does not depend on the count and replays from the CUDA graph)
    k_pois_S_series<<<dim3((m->C + 32 * PSS_CPT - 1) / (32 * PSS_CPT), grid.y), PF_THREADS, 0, m->stream>>>(m->T, s, pa, m->d_pois_ysum[s], out);
    k_pois_chain_list<<<1, 1024, 0, m->stream>>>(m->T, pa, m->d_pois_ysum[s] + pois_stats_doubles((size_t)sp.n_targets) - 1, m->d_pois_list[s]);
    k_pois_S<<<grid, PF_THREADS, 0, m->stream>>>(m->T, s, pa, m->d_pois_ysum[s], out, (const int*)m->d_pois_list[s]);
    m->launches += 2;
  } else
    k_pois_S<<<grid, PF_THREADS, 0, m->stream>>>(m->T, s, pa, m->d_pois_ysum[s], out, (const int*)nullptr);
  m->launches++;
  CK(cudaGetLastError());
  return 0;
}

// evaluate instances [lo,hi) (through idx) of declaration di for all chains; reduced sums land in dst0 (dst1)
static int launch_eval_sum(nimb200_model* m, int mode, int di, int lo, int hi, const int32_t* idx, double* dst0, double* dst1,
                           const nimb200_dep* dp, int slot, int gam) {
  int C = m->C;
  const int gcount = std::max(0, hi - lo);  // size of the selection before it is cut down to this rank's slice
  if (!idx) { lo = std::max(lo, (int)m->shard_lo[di]); hi = std::max(lo, std::min(hi, (int)m->shard_hi[di])); }
  int cwl = cw_log2_for(C), cw = 1 << cwl;
  int count = hi - lo;
  const nimb200_decl& d = m->decl[di];
  const bool norm_family = d.family == NIMB_FAM_NORM_GATHER && m->norm_ok[di] && (mode == 0 || mode == 2);
  if (count <= 0 && !(norm_family && gcount > 0)) {
    CK(cudaMemsetAsync(dst0, 0, sizeof(double) * C, m->stream));
    if (dst1) CK(cudaMemsetAsync(dst1, 0, sizeof(double) * C, m->stream));
    if (m->nranks > 1 && m->sharded[di] && comm_on(m)) {
      if (allreduce_sum(m, dst0, C)) return 1;
      if (dst1 && allreduce_sum(m, dst1, C)) return 1;
    }
    return 0;
  }
  bool shard = m->nranks > 1 && m->sharded[di];
  if (mode == 0 && d.family == NIMB_FAM_BERN_LOGIT_INPROD && m->d_xpad[di] && !idx && (lo % 128) == 0 &&
      (hi == d.n || (hi % 128) == 0)) {
    int nt = fused_bern_logit_tiles(d.n);
    if ((size_t)nt * C > m->partials_cap) return fail("internal: partials scratch too small (fused)");
    fused_bern_logit_launch(m->d_xpad[di], d.n, d.fam_arg[2], m->d_state, d.fam_arg[1], C, m->d_partials, lo / 128,
                            (hi + 127) / 128, m->stream);
    m->launches++;
    CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, nt, 1, C, dst0, dst1));
    m->launches++;
    CK(cudaGetLastError());
    if (shard && comm_on(m)) return allreduce_sum(m, dst0, C);
    return 0;
  }
  if (d.family == NIMB_FAM_NORM_GATHER && m->norm_ok[di] &&
      (mode == 0 || (mode == 2 && dp && dp->param == 0 && (dp->link == NIMB_LINK_ADDITIVE || dp->link == NIMB_LINK_IDENTITY) &&
                     !gam && !(d.fam_arg[3] >= 0 && slot >= d.fam_arg[2] && slot < d.fam_arg[2] + m->norm_vspan[di])))) {
    NormArgs na = norm_args(m, di);
    if (mode == 0 && !idx && gcount == d.n && na.mode >= 0 && !getenv("NIMB200_NO_BYPRODUCT")) {
      // whole-declaration block sum from the data-only group statistics the sweep over the gathered vector keeps: G terms per
      // chain, no pass over the nodes; data-sharded, the statistics were all-reduced once, so every rank forms the full sum
      // itself and nothing is exchanged
      for (size_t s2 = 0; s2 < m->samplers.size(); s2++) {
        const nimb200_sampler& o = m->samplers[s2];
        if (!m->norm_sweep[s2] || !m->norm_gs_valid[s2] || !m->norm_gss_ok[s2] || m->deps[o.dep_off + 1].decl != di) continue;
        if (o.target_slot != d.fam_arg[2] || o.n_targets != m->norm_vspan[di]) continue;
        CK(launch_pdl(k_norm_sum_from_groups, dim3(C), dim3(o.n_targets > 2048 ? 1024 : 256), 0, m->stream, m->T, na, (const double*)m->d_norm_gs[s2],
                      (const double*)m->d_norm_gss[s2], o.n_targets, d.n, dst0));
        m->launches++;
        CK(cudaGetLastError());
        return 0;
      }
    }
    if (mode == 2 && !idx && gcount == d.n && (!shard || comm_on(m)) && na.mode >= 0 && !getenv("NIMB200_NO_BYPRODUCT")) {
      // the sweep over the gathered vector keeps (sum y, count) per group: G terms per chain instead of a pass over the nodes
      for (size_t s2 = 0; s2 < m->samplers.size(); s2++) {
        const nimb200_sampler& o = m->samplers[s2];
        if (!m->norm_sweep[s2] || !m->norm_gs_valid[s2] || m->deps[o.dep_off + 1].decl != di) continue;
        if (o.target_slot != d.fam_arg[2] || o.n_targets != m->norm_vspan[di]) continue;
        CK(launch_pdl(k_norm_conj_from_groups, dim3(C), dim3(o.n_targets > 2048 ? 1024 : 256), 0, m->stream, m->T, na, m->d_norm_gs[s2], o.n_targets, slot, dst0, dst1));
        m->launches++;
        CK(cudaGetLastError());
        return 0;
      }
    }
    int nb = norm_blocks(count);
    int npb = std::max(4 * NF_THREADS, ((std::max(count, 1) + nb - 1) / nb + 4 * NF_THREADS - 1) / (4 * NF_THREADS) * (4 * NF_THREADS));
    nb = std::max(1, (count + npb - 1) / npb);
    if ((size_t)nb * C > m->partials_cap) return fail("internal: partials scratch too small (normal)");
    dim3 grid(nb, (C + NF_CT - 1) / NF_CT);
    // contiguous, 16-byte aligned ranges stream through the TMA ring; index lists and odd offsets use plain loads
    bool tma_ok = !idx && (lo % 4) == 0 && count > 0 && !getenv("NIMB200_NO_TMA");
    if (tma_ok) {
      const int npw = norm_tma_nodes_per_warp(count);
      nb = norm_tma_blocks(count);
      if ((size_t)nb * C > m->partials_cap) return fail("internal: partials scratch too small (normal, staged)");
      grid = dim3(nb, (C + NF_CT - 1) / NF_CT);
      const int ix = na.mode < 0 ? 0 : (na.mode == NIMB_IDX_IARR ? 1 : 2);
#define NIMB_NORM_TMA(M, I)                                                                                        \
  do {                                                                                                             \
    static std::atomic<unsigned long long> attr{0};                                                                \
    ensure_dyn_smem(k_norm_stream_tma<M, I>, NW_SMEM, attr);                                                       \
    k_norm_stream_tma<M, I><<<grid, NF_THREADS, NW_SMEM, m->stream>>>(m->T, na, lo, hi, npw, M == 1 ? slot : -1, m->d_partials); \
  } while (0)
      if (mode == 0) { if (ix == 0) NIMB_NORM_TMA(0, 0); else if (ix == 1) NIMB_NORM_TMA(0, 1); else NIMB_NORM_TMA(0, 2); }
      else { if (ix == 0) NIMB_NORM_TMA(1, 0); else if (ix == 1) NIMB_NORM_TMA(1, 1); else NIMB_NORM_TMA(1, 2); }
#undef NIMB_NORM_TMA
    } else if (mode == 0) k_norm_stream<0><<<grid, NF_THREADS, 0, m->stream>>>(m->T, na, lo, hi, npb, idx, -1, m->d_partials);
    else k_norm_stream<1><<<grid, NF_THREADS, 0, m->stream>>>(m->T, na, lo, hi, npb, idx, slot, m->d_partials);
    m->launches++;
    if (!(shard && comm_on(m))) {
      {  // programmatic dependent launch: the finish kernel is scheduled while the streaming kernel still runs (fused_norm.cuh)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((C * 32 + 127) / 128); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 0; cfg.stream = m->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = getenv("NIMB200_NO_PDL") ? 0 : 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        const double* parts = m->d_partials;
        if (mode == 0) CK(cudaLaunchKernelEx(&cfg, k_norm_reduce_finish<0>, m->T, na, lo, hi, idx, parts, nb, dst0, dst1));
        else CK(cudaLaunchKernelEx(&cfg, k_norm_reduce_finish<1>, m->T, na, lo, hi, idx, parts, nb, dst0, dst1));
      }
    } else {
      CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, nb, 1, C, m->d_normtmp, nullptr));
      m->launches++;
      if (allreduce_sum(m, m->d_normtmp, (size_t)C)) return 1;
      // after the all-reduce the sums cover the whole selection, not only this rank's slice
      if (mode == 0) k_norm_finish<<<(C + 127) / 128, 128, 0, m->stream>>>(m->T, na, gcount, m->d_normtmp, dst0);
      else k_norm_conj_finish<<<(C + 127) / 128, 128, 0, m->stream>>>(m->T, na, gcount, m->d_normtmp, dst0, dst1);
    }
    m->launches++;
    CK(cudaGetLastError());
    return 0;
  }
  if (mode == 0 && d.family == NIMB_FAM_BERN_LOGIT_LIN && m->bern_ok[di] && !idx && C >= 32) {
    BernArgs ba = bern_args(m, di);
    // one wave of equal blocks (3 resident CTAs per SM), each looping over its tiles: with few chains (64 in configs[3]) the
    // grid is a few hundred CTAs and whole-wave quantisation would otherwise idle a third of the SMs
    const int ctiles = (C + 31) / 32;
    const int chunks = std::max(1, std::min(std::min(512, (148 * 3) / std::max(1, ctiles)), (count + 255) / 256));
    int npc = ((count + chunks - 1) / chunks + 31) / 32 * 32;
    if ((size_t)chunks * C > m->partials_cap) return fail("internal: partials scratch too small (bernoulli)");
    dim3 grid((C + 31) / 32, chunks);
    const bool latent = ba.vdata == nullptr, gated = ba.gate_slot >= 0;
    static std::atomic<unsigned long long> a11{0}, a10{0}, a01{0}, a00{0};
    if (latent && gated) {
      ensure_dyn_smem(k_bernlin_sum<true, true>, BF_SUM_SMEM, a11);
      k_bernlin_sum<true, true><<<grid, BF_THREADS, BF_SUM_SMEM, m->stream>>>(m->T, ba, lo, hi, npc, m->d_partials);
    } else if (latent) {
      ensure_dyn_smem(k_bernlin_sum<true, false>, BF_SUM_SMEM, a10);
      k_bernlin_sum<true, false><<<grid, BF_THREADS, BF_SUM_SMEM, m->stream>>>(m->T, ba, lo, hi, npc, m->d_partials);
    } else if (getenv("NIMB200_BERN_GENERAL")) {   // the general kernel on a data declaration (comparison, tests)
      if (gated) {
        ensure_dyn_smem(k_bernlin_sum<false, true>, BF_SUM_SMEM, a01);
        k_bernlin_sum<false, true><<<grid, BF_THREADS, BF_SUM_SMEM, m->stream>>>(m->T, ba, lo, hi, npc, m->d_partials);
      } else {
        ensure_dyn_smem(k_bernlin_sum<false, false>, BF_SUM_SMEM, a00);
        k_bernlin_sum<false, false><<<grid, BF_THREADS, BF_SUM_SMEM, m->stream>>>(m->T, ba, lo, hi, npc, m->d_partials);
      }
    } else if (gated) {
      static std::atomic<unsigned long long> d1{0};
      ensure_dyn_smem(k_bernlin_sum_data<true>, BF_SUM_SMEM, d1);
      k_bernlin_sum_data<true><<<grid, BF_THREADS, BF_SUM_SMEM, m->stream>>>(m->T, ba, lo, hi, npc, m->d_partials);
    } else {
      static std::atomic<unsigned long long> d0{0};
      ensure_dyn_smem(k_bernlin_sum_data<false>, BF_SUM_SMEM, d0);
      k_bernlin_sum_data<false><<<grid, BF_THREADS, BF_SUM_SMEM, m->stream>>>(m->T, ba, lo, hi, npc, m->d_partials);
    }
    m->launches++;
    CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, chunks, 1, C, dst0, dst1));
    m->launches++;
    CK(cudaGetLastError());
    if (shard && comm_on(m)) return allreduce_sum(m, dst0, C);
    return 0;
  }
  if (mode == 0 && d.family == NIMB_FAM_POIS_LOGLIN && m->d_nodeconst[di] && !idx && gcount == d.n) {
    // whole declaration through the group sums: G terms per chain; a pass over the nodes only when the state's b1 is a proposal
    for (size_t s2 = 0; s2 < m->samplers.size(); s2++) {
      if (!pois_S_usable(m, s2) || m->deps[m->samplers[s2].dep_off + 1].decl != di) continue;
      const nimb200_sampler& o = m->samplers[s2];
      PoisArgs pa = pois_args(m, di);
      if (ensure_pois_ysum(m, (int)s2)) return 1;
      const bool b1_proposed = pa.b1_slot >= m->prop_lo && pa.b1_slot < m->prop_hi;
      const double* Suse = m->d_pois_S[s2];
      if (b1_proposed) {
        if (launch_pois_S(m, (int)s2, m->d_pois_Snew[s2])) return 1;
        m->pois_S_pending[s2] = 1;
        Suse = m->d_pois_Snew[s2];
      } else if (!m->pois_S_valid[s2]) {
        if (launch_pois_S(m, (int)s2, m->d_pois_S[s2])) return 1;
        m->pois_S_valid[s2] = 1;
      }
      const int tdecl = m->deps[o.dep_off].decl;
      const bool own = m->nranks > 1 && m->target_sharded[s2];
      const int t_lo = own ? m->shard_lo[tdecl] : 0, t_hi = own ? m->shard_hi[tdecl] : o.n_targets;
      const int chunks = std::max(1, (std::max(0, t_hi - t_lo) + PS_TPC - 1) / PS_TPC);
      if ((size_t)chunks * C > m->partials_cap) break;   // fall through to the node kernel
      k_pois_sum_from_S<<<dim3((C + 31) / 32, chunks), 256, 0, m->stream>>>(m->T, (int)s2, pa, m->d_pois_ysum[s2], Suse, o.n_targets, t_lo, t_hi, m->d_partials);
      m->launches++;
      CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, chunks, 1, C, dst0, dst1));
      m->launches++;
      CK(cudaGetLastError());
      if (own && comm_on(m)) return allreduce_sum(m, dst0, C);
      return 0;
    }
  }
  if (mode == 0 && d.family == NIMB_FAM_POIS_LOGLIN && m->d_nodeconst[di] && !idx && C >= 32) {
    PoisArgs pa = pois_args(m, di);
    const int chunks = family_chunks(C, count, PF_TILE);
    int npc = ((count + chunks - 1) / chunks + PF_TILE - 1) / PF_TILE * PF_TILE;
    if ((size_t)chunks * C > m->partials_cap) return fail("internal: partials scratch too small (poisson)");
    dim3 grid((C + 31) / 32, chunks);
    k_pois_sum<<<grid, PF_THREADS, 0, m->stream>>>(m->T, pa, (const PoisNode*)m->d_nodeconst[di], lo, hi, npc, m->d_partials);
    m->launches++;
    CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, chunks, 1, C, dst0, dst1));
    m->launches++;
    CK(cudaGetLastError());
    if (shard && comm_on(m)) return allreduce_sum(m, dst0, C);
    return 0;
  }
  if (di < (int)m->gam_iid.size() && m->gam_iid[di] && !idx && !shard && lo == 0 && hi == d.n &&
      (mode == 0 || (mode == 2 && gam == 1 && dp && dp->decl == di && (dp->link == NIMB_LINK_IDENTITY || dp->link == NIMB_LINK_MULTIPLICATIVE)))) {
    // i.i.d. gamma layer with chain-level parameters: from (sum log x, sum x), recomputed when a sampler has moved the values
    const int chunks = (d.n + GS_ROWS - 1) / GS_ROWS;
    if ((size_t)chunks * 2 * C <= m->partials_cap) {
      if (!m->gam_stats_valid[di]) {
        CK(launch_pdl(k_gamma_stats, dim3((C + 31) / 32, chunks), dim3(256), 0, m->stream, m->T, di, m->d_partials));
        CK(launch_pdl(k_reduce, dim3((C * 2 * 32 + 127) / 128), dim3(128), 0, m->stream, (const double*)m->d_partials, chunks, 2, C,
                      m->d_gam_stats[di], m->d_gam_stats[di] + C));
        m->launches += 2;
        m->gam_stats_valid[di] = 1;
      }
      nimb200_dep dd{};
      if (dp) dd = *dp;
      if (mode == 0) CK(launch_pdl(k_gamma_from_stats<0>, dim3((C + 3) / 4), dim3(128), 0, m->stream, m->T, di, (const double*)m->d_gam_stats[di], dd, slot, gam, dst0, dst1));
      else CK(launch_pdl(k_gamma_from_stats<2>, dim3((C + 3) / 4), dim3(128), 0, m->stream, m->T, di, (const double*)m->d_gam_stats[di], dd, slot, gam, dst0, dst1));
      m->launches++;
      CK(cudaGetLastError());
      return 0;
    }
  }
  int nt = eval_tiles(C, count);
  int K = mode == 2 ? 2 : 1;
  if ((size_t)nt * K * C > m->partials_cap) return fail("internal: partials scratch too small");
  dim3 grid(nt, (C + cw - 1) / cw);
  nimb200_dep dd{};
  if (dp) dd = *dp;
  size_t shm = sizeof(double) * EVAL_THREADS * K;
  // a single tile (scalar priors, small declarations) is its own sum: the kernel writes the destination directly
  const bool direct = nt == 1 && mode == 0;
  if (mode == 0) CK(launch_pdl(k_eval<0>, grid, dim3(EVAL_THREADS), shm, m->stream, m->T, di, lo, hi, idx, cwl, eval_npt(C, count), direct ? dst0 : m->d_partials, (double*)nullptr, dd, slot, gam));
  else CK(launch_pdl(k_eval<2>, grid, dim3(EVAL_THREADS), shm, m->stream, m->T, di, lo, hi, idx, cwl, eval_npt(C, count), m->d_partials, (double*)nullptr, dd, slot, gam));
  m->launches++;
  if (!direct) {
    CK(launch_pdl(k_reduce, dim3((C * K * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, nt, K, C, dst0, dst1));
    m->launches++;
  }
  CK(cudaGetLastError());
  if (shard && comm_on(m)) {
    if (allreduce_sum(m, dst0, C)) return 1;
    if (dst1 && allreduce_sum(m, dst1, C)) return 1;
  }
  return 0;
}

static int allreduce_sum(nimb200_model* m, double* buf, size_t count) {
  m->allreduces++;
  if (m->xchg_on) {  // one small kernel per piece of at most XCHG_CAP doubles: push to the peers, wait, fixed-order sum
    for (size_t off = 0; off < count; off += XCHG_CAP) {
      const size_t n = std::min(XCHG_CAP, count - off);
      k_peer_allreduce<<<xchg_blocks(n), XCHG_THREADS, 0, m->stream>>>(m->xchg, buf + off, (int)n);
      m->launches++;
    }
    CK(cudaGetLastError());
    return 0;
  }
  int rc = g_nccl.allreduce(buf, buf, count, /*ncclDouble*/ 8, /*ncclSum*/ 0, m->nccl_comm, m->stream);
  if (rc != 0) return fail(std::string("ncclAllReduce failed: ") + (g_nccl.errstr ? g_nccl.errstr(rc) : "?"));
  m->launches++;
  return 0;
}

struct HostSel { int lo, hi; const int32_t* idx; };
static HostSel host_sel(nimb200_model* m, const nimb200_dep& dp, const std::vector<std::vector<int32_t>>& h_iarr) {
  HostSel s{0, 0, nullptr};
  switch (dp.sel) {
    case NIMB_SEL_ALL: s.lo = 0; s.hi = m->decl[dp.decl].n; break;
    case NIMB_SEL_SAME: s.lo = dp.a; s.hi = dp.a + 1; break;
    case NIMB_SEL_STRIDE: s.lo = 0; s.hi = dp.a; break;
    case NIMB_SEL_SEG: s.lo = h_iarr[dp.a][0]; s.hi = h_iarr[dp.a][1]; break;
    case NIMB_SEL_CSR: s.lo = h_iarr[dp.a][0]; s.hi = h_iarr[dp.a][1]; s.idx = m->d_iarr[dp.b]; break;
    default: s.lo = dp.a; s.hi = dp.b; break;
  }
  // NOT cut down to this rank's slice here: launch_eval_sum does that itself and needs the size of the whole selection
  return s;
}

static size_t dep_count_t0(nimb200_model* m, const nimb200_sampler& sp) {
  size_t n = 0;
  for (int e = 0; e < sp.n_dep; e++) {
    HostSel s = host_sel(m, m->deps[sp.dep_off + e], m->h_iarr);
    n += (size_t)std::max(0, s.hi - s.lo);
  }
  return n;
}

static int validate_table(const nimb200_table* t) {
  if (!t) return fail("null table");
  if (t->version != NIMB200_TABLE_VERSION) return fail("node table version mismatch");
  if (t->n_slots < 0 || t->n_decl < 0 || t->n_ops < 0) return fail("negative table size");
  for (int i = 0; i < t->n_decl; i++) {
    const nimb200_decl& d = t->decl[i];
    if (d.dist < NIMB_DNORM || d.dist > NIMB_DLNORM) return fail("unknown distribution id in declaration");
    if (d.n < 0) return fail("negative node count");
    if (d.value_kind == 0) {
      if (d.value_ref < 0 || d.value_ref >= t->n_darr || t->darr_len[d.value_ref] < d.n) return fail("data array too short for declaration");
    } else if (d.value_ref < 0 || d.value_ref + d.n > t->n_slots) return fail("declaration values outside the chain state");
    for (int p = 0; p < d.nparam; p++) {
      if (d.prog_off[p] < 0 || d.prog_off[p] + d.prog_len[p] > t->n_ops) return fail("parameter program outside ops[]");
      int depth = 0, maxd = 0;
      for (int q = 0; q < d.prog_len[p]; q++) {
        int op = t->ops[d.prog_off[p] + q].op;
        if (op >= NIMB_OP_CONST && op <= NIMB_OP_CVARJ) depth++;
        else if (op >= NIMB_OP_ADD && op <= NIMB_OP_POW) depth--;
        else if (op == NIMB_OP_CVAR_IDX || op == NIMB_OP_DATA_IDX) {
          const nimb200_op& o = t->ops[d.prog_off[p] + q];
          if (o.b < 1) return fail("dynamic index op: bad length");
          if (op == NIMB_OP_CVAR_IDX && (o.a < 0 || o.a + o.b > t->n_slots)) return fail("dynamic index op: slots outside the chain state");
          if (op == NIMB_OP_DATA_IDX && (o.a < 0 || o.a >= t->n_darr || t->darr_len[o.a] < o.b)) return fail("dynamic index op: array too short");
        }
        else if (op < NIMB_OP_NEG || op > NIMB_OP_STEP) return fail("unknown op code in parameter program");
        maxd = std::max(maxd, depth);
        if (depth < 1) return fail("malformed parameter program (stack underflow)");
      }
      if (depth != 1) return fail("malformed parameter program");
      if (maxd > 16) return fail("parameter program needs more than 16 stack entries");
    }
  }
  for (int s = 0; s < t->n_samplers; s++) {
    const nimb200_sampler& sp = t->samplers[s];
    if (sp.type < NIMB_SAMPLER_RW || sp.type > NIMB_SAMPLER_CATEGORICAL)
      return fail("unknown sampler type (supported: RW, RW_block, binary, conjugate dgamma/dnorm/dbeta, posterior_predictive, slice, categorical); there is no CPU fallback");
    if (sp.type == NIMB_SAMPLER_SLICE && (sp.tries < 1 || sp.pad < 1 || !(sp.scale > 0))) return fail("slice: sliceWidth, sliceMaxSteps and maxContractions must be positive");
    if (sp.type == NIMB_SAMPLER_RW_BLOCK && (sp.d < 1 || sp.d > NIMB200_MAX_BLOCK_D)) return fail("RW_block dimension out of range");
    if (sp.dep_off < 0 || sp.dep_off + sp.n_dep > t->n_deps || sp.n_dep < 1) return fail("sampler dependency range outside deps[]");
    if (t->deps[sp.dep_off].stage != 0) return fail("first dependency entry of a sampler must be the target's own declaration");
    if (sp.type == NIMB_SAMPLER_CATEGORICAL) {
      const nimb200_decl& cd = t->decl[t->deps[sp.dep_off].decl];
      if (cd.dist != NIMB_DCAT || cd.fam_arg[0] < 1 || cd.fam_arg[0] > NIMB200_MAX_CATEGORIES)
        return fail("categorical: the target must be a dcat node with at most 64 categories");
    }
    if (sp.type == NIMB_SAMPLER_POSTERIOR_PREDICTIVE) {
      const nimb200_decl& pd = t->decl[t->deps[sp.dep_off].decl];
      const bool ok = pd.dist == NIMB_DNORM || pd.dist == NIMB_DGAMMA || pd.dist == NIMB_DBETA || pd.dist == NIMB_DEXP || pd.dist == NIMB_DLNORM ||
                      pd.dist == NIMB_DLOGIS || pd.dist == NIMB_DBINOM || pd.dist == NIMB_DPOIS;
      // dependents (the rest of the predictive network) are allowed: they are simulated by the table samplers that follow this
      // one, and listing them here is what invalidates their cached block sums
      if (!ok) return fail("posterior_predictive: no device variate generator for the target's distribution");
    }
    if (sp.adaptive && sp.adapt_interval < 1) return fail("adaptInterval must be >= 1");
    if (sp.tries < 1 && sp.type == NIMB_SAMPLER_RW_BLOCK) return fail("tries must be >= 1");
  }
  for (int i = 0; i < t->n_monitors; i++)
    if (t->monitor_slots[i] < 0 || t->monitor_slots[i] >= t->n_slots) return fail("monitor slot out of range");
  return 0;
}

extern "C" const char* nimb200_last_error(void) { return g_err.c_str(); }
extern "C" int nimb200_version(void) { return NIMB200_TABLE_VERSION; }

extern "C" int nimb200_model_free(nimb200_model* m) {
  if (!m) return 0;
  cudaSetDevice(m->device);
  if (m->stream) cudaStreamSynchronize(m->stream);
  for (void* p : m->allocs) cudaFree(p);
  if (m->d_samples) cudaFree(m->d_samples);
  if (m->nccl_comm && g_nccl.destroy) g_nccl.destroy(m->nccl_comm);
  for (int r = 0; r < XCHG_MAX_RANKS; r++)
    if (m->xchg_mapped[r]) cudaIpcCloseMemHandle(m->xchg_mapped[r]);
  if (m->xchg_local) cudaFree(m->xchg_local);
  if (m->graph_exec) cudaGraphExecDestroy(m->graph_exec);
  if (m->ev0) cudaEventDestroy(m->ev0);
  if (m->ev1) cudaEventDestroy(m->ev1);
  for (auto e : m->ev_s) cudaEventDestroy(e);
  if (m->stream) cudaStreamDestroy(m->stream);
  delete m;
  return 0;
}

extern "C" int nimb200_model_create(const nimb200_table* t, int32_t nchains, int32_t first_chain, int32_t device,
                                    nimb200_model** out) {
  if (!out) return fail("null output pointer");
  *out = nullptr;
  if (validate_table(t)) return 1;
  if (nchains < 1) return fail("nchains must be >= 1");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
    return fail("no CUDA device available: the B200 back end has no CPU fallback");
  if (device < 0 || device >= ndev) return fail("CUDA device index out of range");
  CK(cudaSetDevice(device));
  nimb200_model* m = new nimb200_model();
  m->device = device; m->C = nchains; m->first_chain = first_chain;
  m->t = *t;
  m->ops.assign(t->ops, t->ops + t->n_ops);
  m->decl.assign(t->decl, t->decl + t->n_decl);
  m->samplers.assign(t->samplers, t->samplers + t->n_samplers);
  m->deps.assign(t->deps, t->deps + t->n_deps);
  m->monitors.assign(t->monitor_slots, t->monitor_slots + t->n_monitors);
  int rc = 0;
#define TRY(x) do { if ((rc = (x)) != 0) { nimb200_model_free(m); return rc; } } while (0)
#define TRYCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fail(std::string("CUDA error ") + cudaGetErrorString(e_)); nimb200_model_free(m); return 1; } } while (0)
  TRYCU(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
  TRYCU(cudaEventCreate(&m->ev0));
  TRYCU(cudaEventCreate(&m->ev1));
  // shared arrays: uploaded once, resident for the life of the model
  m->d_darr.resize(t->n_darr); m->d_iarr.resize(t->n_iarr);
  m->h_iarr.resize(t->n_iarr);
  for (int i = 0; i < t->n_darr; i++) {
    TRY(dev_alloc(m, &m->d_darr[i], (size_t)t->darr_len[i]));
    TRYCU(upload_large(m->d_darr[i], t->darr[i], sizeof(double) * t->darr_len[i], m->stream));
  }
  for (int i = 0; i < t->n_iarr; i++) {
    TRY(dev_alloc(m, &m->d_iarr[i], (size_t)t->iarr_len[i]));
    TRYCU(cudaMemcpyAsync(m->d_iarr[i], t->iarr[i], sizeof(int32_t) * t->iarr_len[i], cudaMemcpyHostToDevice, m->stream));
    m->h_iarr[i].assign(t->iarr[i], t->iarr[i] + t->iarr_len[i]);
  }
  auto up = [&](void** dst, const void* src, size_t bytes) -> int {
    void* q = nullptr;
    CK(cudaMalloc(&q, std::max<size_t>(bytes, 8)));
    m->allocs.push_back(q);
    if (bytes) CK(cudaMemcpyAsync(q, src, bytes, cudaMemcpyHostToDevice, m->stream));
    *dst = q;
    return 0;
  };
  TRY(up(&m->d_ops, m->ops.data(), sizeof(nimb200_op) * m->ops.size()));
  TRY(up(&m->d_decl, m->decl.data(), sizeof(nimb200_decl) * m->decl.size()));
  TRY(up(&m->d_samplers, m->samplers.data(), sizeof(nimb200_sampler) * m->samplers.size()));
  TRY(up(&m->d_deps, m->deps.data(), sizeof(nimb200_dep) * m->deps.size()));
  TRY(up(&m->d_darr_tab, m->d_darr.data(), sizeof(double*) * m->d_darr.size()));
  TRY(up(&m->d_iarr_tab, m->d_iarr.data(), sizeof(int32_t*) * m->d_iarr.size()));
  void* dm = nullptr;
  TRY(up(&dm, m->monitors.data(), sizeof(int32_t) * m->monitors.size()));
  m->d_mon = (int32_t*)dm;
  size_t C = nchains;
  TRY(dev_alloc(m, &m->d_state, (size_t)t->n_slots * C));
  TRY(dev_alloc(m, &m->d_stage, (size_t)t->n_slots * C));
  TRY(dev_alloc(m, &m->d_lpcache, (size_t)t->n_decl * C));
  TRY(dev_alloc(m, &m->d_depcache, (size_t)t->n_deps * C));
  int maxdep = 1;
  for (auto& sp : m->samplers) maxdep = std::max(maxdep, sp.n_dep);
  TRY(dev_alloc(m, &m->d_newsum, (size_t)maxdep * 2 * C));
  TRY(dev_alloc(m, &m->d_tsave, (size_t)NIMB200_MAX_BLOCK_D * C));
  TRY(dev_alloc(m, &m->d_pls, C));
  TRYCU(cudaMemsetAsync(m->d_state, 0, sizeof(double) * t->n_slots * C, m->stream));
  TRYCU(cudaMemsetAsync(m->d_lpcache, 0, sizeof(double) * t->n_decl * C, m->stream));
  // scratch for tile partial sums: sized for the largest declaration
  size_t pc = 1;
  for (auto& d : m->decl) {
    pc = std::max(pc, (size_t)std::max(eval_tiles(nchains, d.n), 592) * 2 * C);  // sub-ranges never need more than 592 tiles
    if (d.family == NIMB_FAM_BERN_LOGIT_INPROD) pc = std::max(pc, (size_t)fused_bern_logit_tiles(d.n) * C);
    if (d.family == NIMB_FAM_POIS_LOGLIN || d.family == NIMB_FAM_BERN_LOGIT_LIN) pc = std::max(pc, (size_t)512 * C);
    if (d.family == NIMB_FAM_NORM_GATHER) pc = std::max(pc, (size_t)(std::max(norm_blocks(d.n), norm_tma_blocks(d.n)) + 1) * 2 * C);
  }
  TRY(dev_alloc(m, &m->d_partials, pc));
  m->partials_cap = pc;
  m->T.ops = (const nimb200_op*)m->d_ops; m->T.decl = (const nimb200_decl*)m->d_decl;
  m->T.deps = (const nimb200_dep*)m->d_deps; m->T.samplers = (const nimb200_sampler*)m->d_samplers;
  m->T.darr = (const double* const*)m->d_darr_tab; m->T.iarr = (const int32_t* const*)m->d_iarr_tab;
  m->T.state = m->d_state; m->T.n_slots = t->n_slots; m->T.C = nchains; m->T.first_chain = first_chain;
  m->shard_lo.assign(m->decl.size(), 0); m->shard_hi.resize(m->decl.size()); m->sharded.assign(m->decl.size(), 0);
  for (size_t i = 0; i < m->decl.size(); i++) m->shard_hi[i] = m->decl[i].n;
  TRY(dev_alloc(m, &m->d_shard_lo, m->decl.size()));
  TRY(dev_alloc(m, &m->d_shard_hi, m->decl.size()));
  TRY(dev_alloc(m, &m->d_sharded, m->decl.size()));
  TRYCU(cudaMemcpyAsync(m->d_shard_lo, m->shard_lo.data(), sizeof(int32_t) * m->decl.size(), cudaMemcpyHostToDevice, m->stream));
  TRYCU(cudaMemcpyAsync(m->d_shard_hi, m->shard_hi.data(), sizeof(int32_t) * m->decl.size(), cudaMemcpyHostToDevice, m->stream));
  TRYCU(cudaMemcpyAsync(m->d_sharded, m->sharded.data(), sizeof(int32_t) * m->decl.size(), cudaMemcpyHostToDevice, m->stream));
  {
    std::vector<int32_t> pc(3 * m->decl.size() + 1, 0);
    for (size_t i = 0; i < m->decl.size(); i++)
      for (int pp = 0; pp < m->decl[i].nparam && pp < 3; pp++) {
        bool cst = true;
        for (int q = 0; q < m->decl[i].prog_len[pp]; q++) {
          int op = m->ops[m->decl[i].prog_off[pp] + q].op;
          if (op == NIMB_OP_DATA || op == NIMB_OP_CVARN || op == NIMB_OP_INPROD || op == NIMB_OP_DATAJ || op == NIMB_OP_CVARJ) cst = false;
        }
        pc[3 * i + pp] = cst ? 1 : 0;
      }
    int32_t* dpc = nullptr;
    TRY(dev_alloc(m, &dpc, pc.size()));
    TRYCU(cudaMemcpyAsync(dpc, pc.data(), sizeof(int32_t) * pc.size(), cudaMemcpyHostToDevice, m->stream));
    TRYCU(cudaStreamSynchronize(m->stream));
    m->T.pconst = dpc;
    m->h_pconst = pc;
    m->gam_iid.assign(m->decl.size(), 0);
    m->gam_stats_valid.assign(m->decl.size(), 0);
    m->d_gam_stats.assign(m->decl.size(), nullptr);
    for (size_t i = 0; i < m->decl.size() && !getenv("NIMB200_NO_GAMMA_STATS"); i++) {
      const nimb200_decl& d = m->decl[i];
      if (d.dist != NIMB_DGAMMA || d.value_kind != 1 || d.nparam < 2 || !pc[3 * i] || !pc[3 * i + 1] || d.n < 256) continue;
      m->gam_iid[i] = 1;
      TRY(dev_alloc(m, &m->d_gam_stats[i], (size_t)2 * m->C));
    }
  }
  TRY(dev_alloc(m, &m->d_ctr, 2));
  TRYCU(cudaMemsetAsync(m->d_ctr, 0, 2 * sizeof(unsigned long long), m->stream));
  TRY(dev_alloc(m, &m->d_ticket, 1));
  { RunParams* rp = nullptr; TRY(dev_alloc(m, &rp, 1)); m->d_runp = rp; }
  TRYCU(cudaMemsetAsync(m->d_ticket, 0, sizeof(unsigned int), m->stream));
  m->T.ctr = m->d_ctr;
  m->T.shard_lo = m->d_shard_lo; m->T.shard_hi = m->d_shard_hi; m->T.sharded = m->d_sharded; m->T.rank = 0; m->T.nranks = 1;
  m->sweep_sharded.assign(m->samplers.size(), 0);
  m->target_sharded.assign(m->samplers.size(), 0);
  // fused family: re-lay X once (row-major, rows padded to a multiple of 4 doubles)
  m->d_xpad.assign(m->decl.size(), nullptr);
  for (size_t i = 0; i < m->decl.size(); i++) {
    nimb200_decl& d = m->decl[i];
    if (d.family != NIMB_FAM_BERN_LOGIT_INPROD) continue;
    int xid = d.fam_arg[0], p = d.fam_arg[2], ld = d.fam_arg[3];
    bool ok = d.value_kind == 0 && xid >= 0 && xid < t->n_darr && p >= 1 && p <= 64 && ld >= p &&
              t->darr_len[xid] >= (int64_t)d.n * ld && d.fam_arg[1] >= 0 && d.fam_arg[1] + p <= t->n_slots;
    for (int q = 0; ok && q < d.n; q++) {  // the fused kernel assumes Bernoulli data in {0, 1}
      double yv = t->darr[d.value_ref][q];
      ok = (yv == 0.0 || yv == 1.0);
    }
    if (!ok) { d.family = NIMB_FAM_GENERIC; continue; }
    size_t len = fused_bern_logit_xpad_len(d.n, p);
    if (len == 0) { d.family = NIMB_FAM_GENERIC; continue; }
    TRY(dev_alloc(m, &m->d_xpad[i], len));
    fused_bern_logit_prepare_x(m->d_darr[xid], m->d_darr[d.value_ref], d.n, p, ld, m->d_xpad[i], m->stream);
  }
  // Bernoulli-logit-linear family: validate
  m->bern_ok.assign(m->decl.size(), 0);
  for (size_t i = 0; i < m->decl.size(); i++) {
    nimb200_decl& d = m->decl[i];
    if (d.family != NIMB_FAM_BERN_LOGIT_LIN) continue;
    int xid = d.fam_arg[0], gs = d.fam_arg[3], gm = d.fam_arg[4], ga = d.fam_arg[5];
    bool ok = d.dist == NIMB_DBINOM && xid >= 0 && xid < t->n_darr && t->darr_len[xid] >= d.n && d.fam_arg[1] >= 0 &&
              d.fam_arg[1] < t->n_slots && d.fam_arg[2] >= 0 && d.fam_arg[2] < t->n_slots && !getenv("NIMB200_NO_FUSED");
    if (ok && gs >= 0) {
      if (gm == NIMB_IDX_IARR) ok = ga >= 0 && ga < t->n_iarr && t->iarr_len[ga] >= d.n;
      if (ok && (gm == NIMB_IDX_DIV || gm == NIMB_IDX_MOD)) ok = ga >= 1;
      int64_t maxg = 0;
      for (int q = 0; ok && q < d.n; q++) {
        int64_t g = gm == NIMB_IDX_I ? q : gm == NIMB_IDX_DIV ? q / ga : gm == NIMB_IDX_MOD ? q % ga : t->iarr[ga][q];
        if (g < 0) ok = false;
        maxg = std::max(maxg, g);
      }
      if (ok) ok = gs + maxg < t->n_slots;
    }
    if (!ok) { d.family = NIMB_FAM_GENERIC; continue; }
    m->bern_ok[i] = 1;
  }
  // streaming normal family: validate
  m->norm_ok.assign(m->decl.size(), 0);
  m->norm_vspan.assign(m->decl.size(), 0);
  TRY(dev_alloc(m, &m->d_normtmp, 2 * C));
  for (size_t i = 0; i < m->decl.size(); i++) {
    nimb200_decl& d = m->decl[i];
    if (d.family != NIMB_FAM_NORM_GATHER) continue;
    int aoff = d.fam_arg[0], alen = d.fam_arg[1], vs = d.fam_arg[2], mode = d.fam_arg[3], arg = d.fam_arg[4];
    bool ok = d.value_kind == 0 && d.dist == NIMB_DNORM && d.nparam == 2 && aoff >= 0 && alen >= 1 && aoff + alen <= t->n_ops &&
              !getenv("NIMB200_NO_FUSED");
    auto chain_level = [&](int off, int len) {
      int depth = 0;
      for (int q = 0; q < len; q++) {
        int op = t->ops[off + q].op;
        if (op == NIMB_OP_DATA || op == NIMB_OP_CVARN || op == NIMB_OP_INPROD || op == NIMB_OP_DATAJ || op == NIMB_OP_CVARJ) return false;
        if (op >= NIMB_OP_CONST && op <= NIMB_OP_CVARJ) depth++;
        else if (op >= NIMB_OP_ADD && op <= NIMB_OP_POW) depth--;
        if (depth < 1 || depth > 16) return false;
      }
      return depth == 1;
    };
    if (ok) ok = chain_level(aoff, alen) && chain_level(d.prog_off[1], d.prog_len[1]);
    int64_t maxg = -1;
    if (ok && mode >= 0) {
      if (mode == NIMB_IDX_IARR) ok = arg >= 0 && arg < t->n_iarr && t->iarr_len[arg] >= d.n;
      if (ok && (mode == NIMB_IDX_DIV || mode == NIMB_IDX_MOD)) ok = arg >= 1;
      for (int q = 0; ok && q < d.n; q++) {
        int64_t g = mode == NIMB_IDX_I ? q : mode == NIMB_IDX_DIV ? q / arg : mode == NIMB_IDX_MOD ? q % arg : t->iarr[arg][q];
        if (g < 0) ok = false;
        maxg = std::max(maxg, g);
      }
      if (ok) ok = vs >= 0 && vs + maxg < t->n_slots;
    }
    if (!ok) { d.family = NIMB_FAM_GENERIC; continue; }
    m->norm_ok[i] = 1;
    m->norm_vspan[i] = (int)(maxg + 1);
  }
  // fused Poisson family: validate, compute the data-only node constants once
  m->d_nodeconst.assign(m->decl.size(), nullptr);
  for (size_t i = 0; i < m->decl.size(); i++) {
    nimb200_decl& d = m->decl[i];
    if (d.family != NIMB_FAM_POIS_LOGLIN) continue;
    int xid = d.fam_arg[0], mode = d.fam_arg[2], arg = d.fam_arg[3];
    bool ok = d.value_kind == 0 && d.dist == NIMB_DPOIS && xid >= 0 && xid < t->n_darr && t->darr_len[xid] >= d.n &&
              d.fam_arg[4] >= 0 && d.fam_arg[4] < t->n_slots && d.fam_arg[5] >= 0 && d.fam_arg[5] < t->n_slots &&
              !getenv("NIMB200_NO_FUSED");
    if (ok && mode == NIMB_IDX_IARR) ok = arg >= 0 && arg < t->n_iarr && t->iarr_len[arg] >= d.n;
    if (ok && (mode == NIMB_IDX_DIV || mode == NIMB_IDX_MOD)) ok = arg >= 1;
    int64_t maxg = 0;
    for (int q = 0; ok && q < d.n; q++) {  // counts must be non-negative integers; lam indices inside the state
      double yv = t->darr[d.value_ref][q];
      ok = yv >= 0.0 && yv == (double)(int64_t)yv && yv < 1e9;
      int64_t g = mode == NIMB_IDX_I ? q : mode == NIMB_IDX_DIV ? q / arg : mode == NIMB_IDX_MOD ? q % arg : t->iarr[arg][q];
      if (g < 0) ok = false;
      maxg = std::max(maxg, g);
    }
    if (ok) ok = d.fam_arg[1] >= 0 && d.fam_arg[1] + maxg < t->n_slots;
    if (!ok) { d.family = NIMB_FAM_GENERIC; continue; }
    PoisNode* pk = nullptr;
    TRY(dev_alloc(m, &pk, (size_t)d.n));
    m->d_nodeconst[i] = pk;
    k_pois_pack<<<(d.n + 255) / 256, 256, 0, m->stream>>>(pois_args(m, (int)i), pk);
  }
  // sampler state + execution shape
  m->sdev.resize(m->samplers.size());
  m->wide.assign(m->samplers.size(), 0);
  m->sweep_deps.assign(m->samplers.size(), 0);
  m->ran.assign(m->samplers.size(), 0);
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    int TT = sp.type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp.n_targets;
    SamplerDev& S = m->sdev[s];
    m->sweep_deps[s] = dep_count_t0(m, sp);
    TRY(dev_alloc(m, &S.scale, (size_t)TT * C));
    TRY(dev_alloc(m, &S.timesRan, (size_t)TT * C));
    TRY(dev_alloc(m, &S.timesAccepted, (size_t)TT * C));
    TRY(dev_alloc(m, &S.timesAdapted, (size_t)TT * C));
    S.propCov = S.chol = S.cholScale = S.empirSamp = nullptr;
    S.hist = nullptr; S.hist_cap = 0; S.hist_K = 0;
    if (sp.type == NIMB_SAMPLER_RW_BLOCK) {
      size_t dd = (size_t)sp.d * sp.d;
      TRY(dev_alloc(m, &S.propCov, dd * C));
      TRY(dev_alloc(m, &S.chol, dd * C));
      TRY(dev_alloc(m, &S.cholScale, dd * C));
      TRY(dev_alloc(m, &S.empirSamp, (size_t)sp.adapt_interval * sp.d * C));
      TRYCU(cudaMemsetAsync(S.empirSamp, 0, sizeof(double) * sp.adapt_interval * sp.d * C, m->stream));
      m->wide[s] = 1;
    } else if (sp.type == NIMB_SAMPLER_CATEGORICAL) {
      if (dep_count_t0(m, sp) > 2048) { nimb200_model_free(m); return fail("categorical sampler over more than 2048 dependents per target is not supported"); }
    } else if (sp.type == NIMB_SAMPLER_SLICE) {
      TRY(dev_alloc(m, &S.propCov, (size_t)TT * C));  // sumJumps
      if (dep_count_t0(m, sp) > 2048) { nimb200_model_free(m); return fail("slice sampler over more than 2048 dependents per target is not supported"); }
    } else if (sp.n_targets == 1 && sp.type != NIMB_SAMPLER_BINARY && sp.type != NIMB_SAMPLER_POSTERIOR_PREDICTIVE &&
               dep_count_t0(m, sp) > 2048) {
      m->wide[s] = 1;
    }
  }
  m->pois_sweep.assign(m->samplers.size(), 0);
  m->d_pois_ysum.assign(m->samplers.size(), nullptr);
  m->pois_ysum_valid.assign(m->samplers.size(), 0);
  m->d_pois_S.assign(m->samplers.size(), nullptr);
  m->d_pois_Snew.assign(m->samplers.size(), nullptr);
  m->d_pois_list.assign(m->samplers.size(), nullptr);
  m->d_pois_chainpar.assign(m->samplers.size(), nullptr);
  m->pois_chain_level.assign(m->samplers.size(), 0);
  m->pois_S_valid.assign(m->samplers.size(), 0);
  m->pois_S_pending.assign(m->samplers.size(), 0);
  TRY(dev_alloc(m, &m->d_accept, (size_t)nchains));
  m->pois_byp_slot.assign(m->samplers.size(), -1);
  m->d_pois_contrib.assign(m->samplers.size(), nullptr);
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    if (sp.type != NIMB_SAMPLER_CONJ_GAMMA || m->wide[s] || sp.n_dep != 2) continue;
    const nimb200_dep& dp = m->deps[sp.dep_off + 1];
    const nimb200_decl& dd = m->decl[dp.decl];
    if (dp.stage == 1 && dd.family == NIMB_FAM_POIS_LOGLIN && m->d_nodeconst[dp.decl] && dp.param == 0 &&
        dp.link == NIMB_LINK_MULTIPLICATIVE && sp.target_slot == dd.fam_arg[1] && m->deps[sp.dep_off].stage == 0) {
      m->pois_sweep[s] = 1;
      TRY(dev_alloc(m, &m->d_pois_ysum[s], pois_stats_doubles((size_t)sp.n_targets)));
    }
  }
  m->bern_sweep.assign(m->samplers.size(), 0);
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    if (sp.type != NIMB_SAMPLER_BINARY || m->wide[s] || sp.n_dep != 2) continue;
    const nimb200_dep &d0 = m->deps[sp.dep_off], &d1 = m->deps[sp.dep_off + 1];
    const nimb200_decl &z = m->decl[d0.decl], &y = m->decl[d1.decl];
    if (d0.stage == 0 && d1.stage == 1 && z.family == NIMB_FAM_BERN_LOGIT_LIN && m->bern_ok[d0.decl] && z.value_kind == 1 &&
        z.fam_arg[3] < 0 && (d0.sel == NIMB_SEL_SAME || d0.sel == NIMB_SEL_RANGE) &&
        z.value_ref + (d0.sel == NIMB_SEL_SAME ? d0.a : d0.a) == sp.target_slot && y.family == NIMB_FAM_BERN_LOGIT_LIN &&
        m->bern_ok[d1.decl] && y.value_kind == 0 && y.fam_arg[3] == sp.target_slot)
      m->bern_sweep[s] = 1;
  }
  m->norm_sweep.assign(m->samplers.size(), 0);
  m->d_norm_gs.assign(m->samplers.size(), nullptr);
  m->norm_gs_valid.assign(m->samplers.size(), 0);
  m->d_norm_gss.assign(m->samplers.size(), nullptr);
  m->norm_gss_ok.assign(m->samplers.size(), 0);
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    if (sp.type != NIMB_SAMPLER_CONJ_NORMAL || m->wide[s] || sp.n_dep != 2) continue;
    const nimb200_dep& dp = m->deps[sp.dep_off + 1];
    const nimb200_decl& dd = m->decl[dp.decl];
    if (!(dp.stage == 1 && dd.family == NIMB_FAM_NORM_GATHER && m->norm_ok[dp.decl] && dd.fam_arg[3] >= 0 && dp.param == 0 &&
          (dp.link == NIMB_LINK_ADDITIVE || dp.link == NIMB_LINK_IDENTITY) && sp.target_slot == dd.fam_arg[2] &&
          m->deps[sp.dep_off].stage == 0)) continue;
    bool a_reads_target = false;  // A must not read the gathered vector itself
    for (int q = 0; q < dd.fam_arg[1]; q++) {
      const nimb200_op& o = m->ops[dd.fam_arg[0] + q];
      if (o.op == NIMB_OP_CVAR && o.a >= sp.target_slot && o.a < sp.target_slot + sp.n_targets) a_reads_target = true;
    }
    if (!a_reads_target) {
      m->norm_sweep[s] = 1;
      TRY(dev_alloc(m, &m->d_norm_gs[s], (size_t)sp.n_targets * 2));
      TRY(dev_alloc(m, &m->d_norm_gss[s], (size_t)sp.n_targets));
    }
  }
  // Cache slots: the block sum of the CURRENT state over (declaration, selection).  Entries of wide samplers whose selection
  // does not depend on the target share a slot, so b1's step can reuse what b0's step just committed instead of re-evaluating
  // the dependents.  A sampler invalidates every slot on a declaration it touches, except -- for a wide MH sampler -- the slots
  // of its own entries, which its accept/reject step keeps equal to the state it leaves behind.
  m->dep_slot.assign(m->deps.size(), 0);
  {
    std::vector<std::array<int, 4>> keys;
    for (size_t g = 0; g < m->deps.size(); g++) {
      const nimb200_dep& dp = m->deps[g];
      std::array<int, 4> key = {dp.decl, dp.sel, dp.a, dp.b};
      bool shareable = dp.sel == NIMB_SEL_ALL || dp.sel == NIMB_SEL_RANGE;
      int slot = -1;
      if (shareable)
        for (size_t q = 0; q < keys.size(); q++)
          if (keys[q] == key) { slot = (int)q; break; }
      if (slot < 0) { slot = (int)keys.size(); keys.push_back(shareable ? key : std::array<int, 4>{-1, (int)g, 0, 0}); }
      m->dep_slot[g] = slot;
    }
    m->dep_valid.assign(keys.size(), 0);
    void* ds = nullptr;
    TRY(up(&ds, m->dep_slot.data(), sizeof(int32_t) * m->dep_slot.size()));
    m->d_dep_slot = (int32_t*)ds;
  }
  // A Poisson conjugate sweep whose targets together cover every node of the dependent declaration can hand the next MH step on
  // that declaration its current-state block sum (fused_pois.cuh): needs a cache slot keyed on the whole declaration, counts
  // small enough for the closed form (<= PF_YMAX) and no data sharding at run time.
  for (size_t s = 0; s < m->samplers.size(); s++) {
    if (!m->pois_sweep[s] || getenv("NIMB200_NO_BYPRODUCT")) continue;
    const nimb200_sampler& sp = m->samplers[s];
    const nimb200_dep& dp = m->deps[sp.dep_off + 1];
    const nimb200_decl& dd = m->decl[dp.decl];
    bool small = true;
    for (int q = 0; q < dd.n && small; q++) small = t->darr[dd.value_ref][q] <= PF_YMAX;
    // every node belongs to exactly one target?  (index map onto the targets: each node reads one lam element inside the block)
    bool covers = dd.fam_arg[1] == sp.target_slot;
    int64_t hit = 0;
    {
      const int mode = dd.fam_arg[2], arg = dd.fam_arg[3];
      for (int q = 0; q < dd.n && covers; q++) {
        int64_t g = mode == NIMB_IDX_I ? q : mode == NIMB_IDX_DIV ? q / arg : mode == NIMB_IDX_MOD ? q % arg : t->iarr[arg][q];
        covers = g >= 0 && g < sp.n_targets;
        hit++;
      }
    }
    if (!small || !covers || hit != dd.n) continue;
    for (size_t g = 0; g < m->deps.size(); g++) {
      const nimb200_dep& o = m->deps[g];
      bool whole = o.decl == dp.decl && (o.sel == NIMB_SEL_ALL || (o.sel == NIMB_SEL_RANGE && o.a == 0 && o.b == dd.n));
      if (whole && ((int)g < sp.dep_off || (int)g >= sp.dep_off + sp.n_dep)) { m->pois_byp_slot[s] = m->dep_slot[g]; break; }
    }
    if (m->pois_byp_slot[s] >= 0) TRY(dev_alloc(m, &m->d_pois_contrib[s], (size_t)sp.n_targets * C));
    // same eligibility (targets cover the declaration, small counts) for the group-sum path; two [targets][chains] arrays
    if (m->pois_byp_slot[s] >= 0 && !getenv("NIMB200_NO_GROUPSUMS") && (size_t)sp.n_targets * C * 16 < ((size_t)8 << 30)) {
      m->pois_series = !getenv("NIMB200_NO_SERIES");
      TRY(dev_alloc(m, &m->d_pois_S[s], (size_t)sp.n_targets * C));
      TRY(dev_alloc(m, &m->d_pois_Snew[s], (size_t)sp.n_targets * C));
      TRY(dev_alloc(m, &m->d_pois_list[s], (size_t)C + 1));
      TRY(dev_alloc(m, &m->d_pois_chainpar[s], (size_t)3 * C));
      const int tdecl = m->deps[sp.dep_off].decl;
      m->pois_chain_level[s] = m->decl[tdecl].nparam >= 2 && m->h_pconst[3 * tdecl] && m->h_pconst[3 * tdecl + 1];
    }
  }
  // The binary sweep of the occupancy family evaluates, for the state it leaves behind, every node of the sites' own declaration and
  // of the visits' declaration: when its targets cover both declarations it refreshes their cached block sums (fused_bern.cuh).
  m->bern_byp_slot.assign(m->samplers.size(), std::array<int, 2>{-1, -1});
  m->d_bern_contrib.assign(m->samplers.size(), nullptr);
  for (size_t s = 0; s < m->samplers.size(); s++) {
    if (!m->bern_sweep[s] || getenv("NIMB200_NO_BYPRODUCT")) continue;
    const nimb200_sampler& sp = m->samplers[s];
    const nimb200_dep &d0 = m->deps[sp.dep_off], &d1 = m->deps[sp.dep_off + 1];
    const nimb200_decl &z = m->decl[d0.decl], &y = m->decl[d1.decl];
    bool covers = sp.n_targets == z.n && z.value_ref == sp.target_slot;
    {
      const int mode = y.fam_arg[4], arg = y.fam_arg[5];
      for (int q = 0; q < y.n && covers; q++) {
        int64_t g = mode == NIMB_IDX_I ? q : mode == NIMB_IDX_DIV ? q / arg : mode == NIMB_IDX_MOD ? q % arg : t->iarr[arg][q];
        covers = g >= 0 && g < sp.n_targets;
      }
    }
    if (!covers) continue;
    std::array<int, 2> slots{-1, -1};
    for (size_t g = 0; g < m->deps.size(); g++) {
      if ((int)g >= sp.dep_off && (int)g < sp.dep_off + sp.n_dep) continue;
      const nimb200_dep& o = m->deps[g];
      for (int k = 0; k < 2; k++) {
        const int decl = k == 0 ? d0.decl : d1.decl, n = k == 0 ? z.n : y.n;
        if (slots[k] < 0 && o.decl == decl && (o.sel == NIMB_SEL_ALL || (o.sel == NIMB_SEL_RANGE && o.a == 0 && o.b == n))) slots[k] = m->dep_slot[g];
      }
    }
    if (slots[0] < 0 || slots[1] < 0) continue;
    if ((size_t)((sp.n_targets + CS_ROWS - 1) / CS_ROWS) * C > m->partials_cap) continue;  // column-sum partials live in d_partials
    m->bern_byp_slot[s] = slots;
    TRY(dev_alloc(m, &m->d_bern_contrib[s], (size_t)2 * sp.n_targets * C));
  }
  {  // whole-iteration kernel: every sampler a plain scalar sweep with little work
    SmallPlan P{};
    bool ok = !getenv("NIMB200_NO_SMALL") && m->samplers.size() <= 24 && (size_t)t->n_monitors * C <= 65536;
    for (size_t s = 0; s < m->samplers.size() && ok; s++) {
      const nimb200_sampler& sp = m->samplers[s];
      const bool plain = sp.type == NIMB_SAMPLER_RW || sp.type == NIMB_SAMPLER_BINARY || sp.type == NIMB_SAMPLER_CONJ_GAMMA ||
                         sp.type == NIMB_SAMPLER_CONJ_NORMAL || sp.type == NIMB_SAMPLER_CONJ_BETA || sp.type == NIMB_SAMPLER_POSTERIOR_PREDICTIVE;
      if (!plain || m->wide[s] || m->pois_sweep[s] || m->bern_sweep[s] || m->norm_sweep[s]) { ok = false; break; }
      const size_t pairs = (size_t)sp.n_targets * C;
      const bool wpp = pairs <= 65536 && m->sweep_deps[s] >= 8;      // the same choice run_sampler makes
      if (pairs * (wpp ? 32 : 1) > 8192 || m->sweep_deps[s] > 4096) { ok = false; break; }
      P.sid[P.n] = (int)s; P.lanes[P.n] = wpp ? 32 : 1; P.n++;
    }
    if (ok && P.n > 0) {
      m->small = P;
      TRY(dev_alloc(m, &m->d_sdev, m->samplers.size()));
      TRYCU(cudaMemcpyAsync(m->d_sdev, m->sdev.data(), sizeof(SamplerDev) * m->sdev.size(), cudaMemcpyHostToDevice, m->stream));
    }
  }
  m->invalidates.resize(m->samplers.size());
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    std::set<int> touched, own, inval;
    bool keeps_own = m->wide[s] && sp.type != NIMB_SAMPLER_CONJ_GAMMA && sp.type != NIMB_SAMPLER_CONJ_NORMAL && sp.type != NIMB_SAMPLER_CONJ_BETA;
    for (int e = 0; e < sp.n_dep; e++) {
      touched.insert(m->deps[sp.dep_off + e].decl);
      if (keeps_own) own.insert(m->dep_slot[sp.dep_off + e]);
    }
    for (size_t g = 0; g < m->deps.size(); g++)
      if (touched.count(m->deps[g].decl) && !own.count(m->dep_slot[g])) inval.insert(m->dep_slot[g]);
    m->invalidates[s].assign(inval.begin(), inval.end());
  }
  TRYCU(cudaStreamSynchronize(m->stream));
  *out = m;
  return 0;
}

extern "C" int nimb200_set_values(nimb200_model* m, const double* v) {
  if (!m || !v) return fail("null argument");
  CK(cudaSetDevice(m->device));
  size_t n = (size_t)m->t.n_slots * m->C;
  if (n == 0) return 0;
  CK(cudaMemcpyAsync(m->d_stage, v, sizeof(double) * n, cudaMemcpyHostToDevice, m->stream));
  k_transpose<<<(unsigned)((n + 255) / 256), 256, 0, m->stream>>>(m->d_stage, m->d_state, m->C, m->t.n_slots);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(m->stream));
  std::fill(m->dep_valid.begin(), m->dep_valid.end(), 0);
  std::fill(m->lp_valid.begin(), m->lp_valid.end(), 0);
  std::fill(m->gam_stats_valid.begin(), m->gam_stats_valid.end(), 0);
  std::fill(m->pois_S_valid.begin(), m->pois_S_valid.end(), 0);
  return 0;
}

extern "C" int nimb200_get_values(nimb200_model* m, double* v) {
  if (!m || !v) return fail("null argument");
  CK(cudaSetDevice(m->device));
  size_t n = (size_t)m->t.n_slots * m->C;
  if (n == 0) return 0;
  k_transpose<<<(unsigned)((n + 255) / 256), 256, 0, m->stream>>>(m->d_state, m->d_stage, m->t.n_slots, m->C);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(v, m->d_stage, sizeof(double) * n, cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  return 0;
}

static int check_range(nimb200_model* m, int di, int lo, int hi) {
  if (!m) return fail("null model");
  if (di < 0 || di >= (int)m->decl.size()) return fail("declaration index out of range");
  if (lo < 0 || hi > m->decl[di].n || lo > hi) return fail("node range outside the declaration");
  return 0;
}

extern "C" int nimb200_calculate(nimb200_model* m, int32_t di, int32_t lo, int32_t hi, double* out) {
  if (check_range(m, di, lo, hi)) return 1;
  CK(cudaSetDevice(m->device));
  bool whole = lo == 0 && hi == m->decl[di].n;
  double* dst = whole ? m->d_lpcache + (size_t)di * m->C : m->d_pls;
  if (launch_eval_sum(m, 0, di, lo, hi, nullptr, dst, nullptr, nullptr, -1, 0)) return 1;
  CK(cudaMemcpyAsync(out, dst, sizeof(double) * m->C, cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  if (whole) { if (m->lp_valid.size() != m->decl.size()) m->lp_valid.assign(m->decl.size(), 0); m->lp_valid[di] = 1; }
  return 0;
}

extern "C" int nimb200_calculate_nodes(nimb200_model* m, int32_t di, int32_t lo, int32_t hi, double* out) {
  if (check_range(m, di, lo, hi)) return 1;
  CK(cudaSetDevice(m->device));
  size_t n = (size_t)(hi - lo) * m->C;
  if (n == 0) return 0;
  double* d_out = nullptr;
  CK(cudaMalloc(&d_out, sizeof(double) * n));
  int cwl = cw_log2_for(m->C), cw = 1 << cwl;
  dim3 grid(eval_tiles(m->C, hi - lo), (m->C + cw - 1) / cw);
  nimb200_dep dd{};
  k_eval<1><<<grid, EVAL_THREADS, 0, m->stream>>>(m->T, di, lo, hi, nullptr, cwl, eval_npt(m->C, hi - lo), nullptr, d_out, dd, -1, 0);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, sizeof(double) * n, cudaMemcpyDeviceToHost, m->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
  cudaFree(d_out);
  if (e != cudaSuccess) return fail(std::string("CUDA error ") + cudaGetErrorString(e));
  return 0;
}

__global__ void k_diff_update(double* cache, const double* nw, double* out, int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  out[c] = nw[c] - cache[c];
  cache[c] = nw[c];
}

extern "C" int nimb200_calculateDiff(nimb200_model* m, int32_t di, double* out) {
  if (check_range(m, di, 0, 0)) return 1;
  CK(cudaSetDevice(m->device));
  if (launch_eval_sum(m, 0, di, 0, m->decl[di].n, nullptr, m->d_newsum, nullptr, nullptr, -1, 0)) return 1;
  k_diff_update<<<(m->C + 127) / 128, 128, 0, m->stream>>>(m->d_lpcache + (size_t)di * m->C, m->d_newsum, m->d_pls, m->C);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, m->d_pls, sizeof(double) * m->C, cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  return 0;
}

// bench.py: average device time of `reps` back-to-back launches of the block-sum kernel of declaration di
// (the kernel calculate() uses), CUDA events on the launching stream.
extern "C" int nimb200_time_calculate(nimb200_model* m, int32_t di, int32_t reps, double* ms_avg) {
  if (check_range(m, di, 0, 0)) return 1;
  if (reps < 1 || !ms_avg) return fail("bad argument");
  CK(cudaSetDevice(m->device));
  double* dst = m->d_newsum;
  if (launch_eval_sum(m, 0, di, 0, m->decl[di].n, nullptr, dst, nullptr, nullptr, -1, 0)) return 1;  // warm
  CK(cudaStreamSynchronize(m->stream));
  const nimb200_decl& d = m->decl[di];
  bool fused = d.family == NIMB_FAM_BERN_LOGIT_INPROD && m->d_xpad[di];
  int C = m->C, cwl = cw_log2_for(C), cw = 1 << cwl;
  bool family = d.family == NIMB_FAM_NORM_GATHER || d.family == NIMB_FAM_POIS_LOGLIN || d.family == NIMB_FAM_BERN_LOGIT_LIN;
  CK(cudaEventRecord(m->ev0, m->stream));
  for (int r = 0; r < reps; r++) {
    if (family) {  // streaming families: the whole block-sum evaluation (main kernel + the small reduce / finish kernels)
      if (launch_eval_sum(m, 0, di, 0, d.n, nullptr, dst, nullptr, nullptr, -1, 0)) return 1;
    } else if (fused)
      fused_bern_logit_launch(m->d_xpad[di], d.n, d.fam_arg[2], m->d_state, d.fam_arg[1], C, m->d_partials, 0, (d.n + 127) / 128, m->stream);
    else {
      dim3 grid(eval_tiles(C, d.n), (C + cw - 1) / cw);
      nimb200_dep dd{};
      k_eval<0><<<grid, EVAL_THREADS, sizeof(double) * EVAL_THREADS, m->stream>>>(m->T, di, 0, d.n, nullptr, cwl, eval_npt(C, d.n), m->d_partials, nullptr, dd, -1, 0);
    }
  }
  CK(cudaEventRecord(m->ev1, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  CK(cudaGetLastError());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, m->ev0, m->ev1));
  *ms_avg = ms / reps;
  return 0;
}

// average device milliseconds of one all-reduce of `count` doubles on the model's run stream, through whichever transport the
// data-sharded model was given (every rank of the group must make the same call)
extern "C" int nimb200_time_allreduce(nimb200_model* m, int32_t count, int32_t reps, double* ms_avg) {
  if (!m || !ms_avg || reps < 1 || count < 1) return fail("bad argument");
  if (!comm_on(m)) return fail("the model has no all-reduce transport (not data-sharded)");
  if ((size_t)count > (size_t)2 * m->C) return fail("count exceeds the scratch vector (2 x chains)");
  CK(cudaSetDevice(m->device));
  CK(cudaMemsetAsync(m->d_newsum, 0, sizeof(double) * count, m->stream));
  if (allreduce_sum(m, m->d_newsum, count)) return 1;  // warm
  CK(cudaStreamSynchronize(m->stream));
  CK(cudaEventRecord(m->ev0, m->stream));
  for (int r = 0; r < reps; r++)
    if (allreduce_sum(m, m->d_newsum, count)) return 1;
  CK(cudaEventRecord(m->ev1, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  CK(cudaGetLastError());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, m->ev0, m->ev1));
  *ms_avg = ms / reps;
  return 0;
}

extern "C" int nimb200_getLogProb(nimb200_model* m, int32_t di, double* out) {
  if (check_range(m, di, 0, 0)) return 1;
  CK(cudaSetDevice(m->device));
  CK(cudaMemcpyAsync(out, m->d_lpcache + (size_t)di * m->C, sizeof(double) * m->C, cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  return 0;
}

// ------------------------------------------------------------------------------------------------ data-node sharding
extern "C" int nimb200_nccl_unique_id(void* id128) {
  if (!id128) return fail("null argument");
  if (load_nccl()) return fail("NCCL library (libnccl.so.2) could not be loaded");
  NcclId id;
  int rc = g_nccl.getid(&id);
  if (rc != 0) return fail("ncclGetUniqueId failed");
  memcpy(id128, &id, sizeof id);
  return 0;
}

// Every rank of a data-sharded group calls this on its own model (same table, same chains, same seed) before the
// first run.  Data declarations with at least min_nodes instances are split into contiguous, 128-aligned slices;
// id128 == NULL keeps the slices but creates no communicator (block sums are then this rank's partial sums: used by
// tests that add the parts up themselves).
static int set_shards(nimb200_model* m, int32_t nranks, int32_t rank, int32_t min_nodes) {
  if (!m) return fail("null model");
  if (nranks < 1 || rank < 0 || rank >= nranks) return fail("bad rank / nranks");
  CK(cudaSetDevice(m->device));
  m->rank = rank; m->nranks = nranks;
  std::fill(m->norm_gs_valid.begin(), m->norm_gs_valid.end(), 0);
  std::fill(m->pois_ysum_valid.begin(), m->pois_ysum_valid.end(), 0);
  for (size_t i = 0; i < m->decl.size(); i++) {
    const nimb200_decl& d = m->decl[i];
    bool sh = nranks > 1 && d.value_kind == 0 && d.n >= min_nodes;
    m->sharded[i] = sh ? 1 : 0;
    if (sh) {
      int64_t per = ((int64_t)d.n + nranks - 1) / nranks;
      per = (per + 127) / 128 * 128;
      m->shard_lo[i] = (int32_t)std::min<int64_t>(d.n, per * rank);
      m->shard_hi[i] = (int32_t)std::min<int64_t>(d.n, per * (rank + 1));
    } else { m->shard_lo[i] = 0; m->shard_hi[i] = d.n; }
  }
  // Group-aligned shards: a conjugate gamma-Poisson sweep whose targets own contiguous runs of the sharded declaration's nodes
  // (sorted groups) gets its shard boundaries moved onto target boundaries.  Every target's nodes then sit on ONE rank, that
  // rank alone draws the target (no exchange of [targets][chains] rate sums, 655 MB per sweep for the GLMM at 8192 chains), and
  // the target declaration itself becomes sharded by ownership: its block sums are formed from each rank's own targets and
  // all-reduced like any other sharded declaration's.  Not done when a target is monitored (the other ranks' copies are stale).
  std::fill(m->target_sharded.begin(), m->target_sharded.end(), 0);
  for (size_t s = 0; s < m->samplers.size() && nranks > 1 && !getenv("NIMB200_NO_GROUP_SHARDS"); s++) {
    if (!m->pois_sweep[s]) continue;
    const nimb200_sampler& sp = m->samplers[s];
    const nimb200_dep &tp = m->deps[sp.dep_off], &dp = m->deps[sp.dep_off + 1];
    const nimb200_decl &td = m->decl[tp.decl], &dd = m->decl[dp.decl];
    if (!m->sharded[dp.decl] || td.value_kind != 1 || tp.sel != NIMB_SEL_SAME || tp.a != 0 || sp.n_targets != td.n) continue;
    bool claimed = false;   // one owner layout per declaration
    for (size_t s2 = 0; s2 < s; s2++) claimed |= m->target_sharded[s2] && m->deps[m->samplers[s2].dep_off + 1].decl == dp.decl;
    bool monitored = false;
    for (int32_t slot : m->monitors) monitored |= slot >= sp.target_slot && slot < sp.target_slot + sp.n_targets;
    if (claimed || monitored) continue;
    std::vector<int64_t> off((size_t)sp.n_targets + 1);
    if (dp.sel == NIMB_SEL_SAME) for (int t = 0; t <= sp.n_targets; t++) off[t] = (int64_t)dp.a + t;
    else if (dp.sel == NIMB_SEL_STRIDE) for (int t = 0; t <= sp.n_targets; t++) off[t] = (int64_t)dp.a * t;
    else if (dp.sel == NIMB_SEL_SEG) for (int t = 0; t <= sp.n_targets; t++) off[t] = m->h_iarr[dp.a][t];
    else continue;
    auto first_target_at = [&](int r) {  // first target whose nodes start at or after rank r's even share
      const int64_t want = (int64_t)dd.n * r / nranks;
      return (int)(std::lower_bound(off.begin(), off.end() - 1, want) - off.begin());
    };
    const int t_lo = rank == 0 ? 0 : first_target_at(rank), t_hi = rank == nranks - 1 ? sp.n_targets : first_target_at(rank + 1);
    m->shard_lo[dp.decl] = rank == 0 ? 0 : (int32_t)off[t_lo];
    m->shard_hi[dp.decl] = rank == nranks - 1 ? dd.n : (int32_t)off[t_hi];
    m->sharded[tp.decl] = 1;
    m->shard_lo[tp.decl] = t_lo;
    m->shard_hi[tp.decl] = t_hi;
    m->target_sharded[s] = 1;
  }
  CK(cudaMemcpyAsync(m->d_shard_lo, m->shard_lo.data(), sizeof(int32_t) * m->decl.size(), cudaMemcpyHostToDevice, m->stream));
  CK(cudaMemcpyAsync(m->d_shard_hi, m->shard_hi.data(), sizeof(int32_t) * m->decl.size(), cudaMemcpyHostToDevice, m->stream));
  CK(cudaMemcpyAsync(m->d_sharded, m->sharded.data(), sizeof(int32_t) * m->decl.size(), cudaMemcpyHostToDevice, m->stream));
  m->T.rank = rank; m->T.nranks = nranks;
  size_t need = 0;
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    bool touch = false;
    for (int e = 0; e < sp.n_dep; e++) touch |= m->sharded[m->deps[sp.dep_off + e].decl] != 0;
    m->sweep_sharded[s] = touch ? 1 : 0;
    if (touch && !m->wide[s]) need = std::max(need, (size_t)sp.n_targets * 3 * m->C);
  }
  if (need > m->sweepbuf_cap) {
    if (dev_alloc(m, &m->d_sweepbuf, need)) return 1;
    m->sweepbuf_cap = need;
  }
  CK(cudaStreamSynchronize(m->stream));
  std::fill(m->dep_valid.begin(), m->dep_valid.end(), 0);
  std::fill(m->lp_valid.begin(), m->lp_valid.end(), 0);
  std::fill(m->gam_stats_valid.begin(), m->gam_stats_valid.end(), 0);
  std::fill(m->pois_S_valid.begin(), m->pois_S_valid.end(), 0);
  if (m->graph_exec) { cudaGraphExecDestroy(m->graph_exec); m->graph_exec = nullptr; }
  return 0;
}

extern "C" int nimb200_enable_data_sharding(nimb200_model* m, int32_t nranks, int32_t rank, const void* id128, int32_t min_nodes) {
  if (set_shards(m, nranks, rank, min_nodes)) return 1;
  if (nranks > 1 && id128) {
    if (load_nccl()) return fail("NCCL library (libnccl.so.2) could not be loaded");
    NcclId id;
    memcpy(&id, id128, sizeof id);
    int rc = g_nccl.init(&m->nccl_comm, nranks, id, rank);
    if (rc != 0) return fail(std::string("ncclCommInitRank failed: ") + (g_nccl.errstr ? g_nccl.errstr(rc) : "?"));
  }
  return 0;
}

// ---- peer exchange: the all-reduce of the data-sharded mode as a kernel of ours over NVLink peer memory (peer_xchg.cuh)
namespace {
struct PeerHandle {          // what nimb200_peer_export writes into the caller's 128 bytes
  uint64_t magic;            // 'NIMBXCHG'
  uint64_t pid;              // exporting process: inside one process peers use the raw pointer, across processes CUDA IPC
  uint64_t ptr;              // device address in the exporting process
  int32_t device, nranks;
  cudaIpcMemHandle_t ipc;    // 64 bytes
};
static_assert(sizeof(PeerHandle) <= NIMB200_PEER_HANDLE_BYTES, "peer handle does not fit");
constexpr uint64_t XCHG_MAGIC = 0x4e494d4258434847ull;
}  // namespace

extern "C" int nimb200_peer_export(nimb200_model* m, int32_t nranks, void* handle128) {
  if (!m || !handle128) return fail("null argument");
  if (nranks < 2 || nranks > XCHG_MAX_RANKS) return fail("peer exchange supports 2 to 8 ranks (one NVSwitch box)");
  CK(cudaSetDevice(m->device));
  if (m->xchg_on) return fail("peer exchange is already attached");
  if (m->xchg_local) { cudaFree(m->xchg_local); m->xchg_local = nullptr; }
  CK(cudaMalloc(&m->xchg_local, xchg_bytes(nranks)));
  CK(cudaMemset(m->xchg_local, 0, xchg_bytes(nranks)));   // flags and sequence number start at zero on every rank ...
  CK(cudaDeviceSynchronize());                            // ... before any peer can learn this address
  PeerHandle h{};
  h.magic = XCHG_MAGIC; h.pid = (uint64_t)getpid(); h.ptr = (uint64_t)(uintptr_t)m->xchg_local; h.device = m->device; h.nranks = nranks;
  CK(cudaIpcGetMemHandle(&h.ipc, m->xchg_local));
  memset(handle128, 0, NIMB200_PEER_HANDLE_BYTES);
  memcpy(handle128, &h, sizeof h);
  return 0;
}

extern "C" int nimb200_enable_data_sharding_peer(nimb200_model* m, int32_t nranks, int32_t rank, const void* handles,
                                                 int32_t min_nodes) {
  if (!m || !handles) return fail("null argument");
  if (nranks < 2 || nranks > XCHG_MAX_RANKS) return fail("peer exchange supports 2 to 8 ranks (one NVSwitch box)");
  if (rank < 0 || rank >= nranks) return fail("bad rank / nranks");
  if (!m->xchg_local) return fail("call nimb200_peer_export on this model first");
  if (m->nccl_comm) return fail("this model already all-reduces over NCCL");
  if (set_shards(m, nranks, rank, min_nodes)) return 1;
  const char* hb = (const char*)handles;
  PeerXchg X{};
  X.nranks = nranks; X.rank = rank;
  for (int r = 0; r < nranks; r++) {
    PeerHandle h;
    memcpy(&h, hb + (size_t)r * NIMB200_PEER_HANDLE_BYTES, sizeof h);
    if (h.magic != XCHG_MAGIC || h.nranks != nranks) return fail("peer handle " + std::to_string(r) + " is not a handle of this exchange");
    void* base = nullptr;
    if (r == rank) {
      if ((void*)(uintptr_t)h.ptr != m->xchg_local) return fail("handle[rank] is not this model's own export");
      base = m->xchg_local;
    } else if (h.pid == (uint64_t)getpid() && h.device == m->device) {
      base = (void*)(uintptr_t)h.ptr;          // same process, same device (two ranks sharing a GPU: tests on a one-GPU box)
    } else if (h.pid == (uint64_t)getpid()) {  // same process, another device: plain peer access
      int can = 0;
      CK(cudaDeviceCanAccessPeer(&can, m->device, h.device));
      if (!can) return fail("device " + std::to_string(m->device) + " cannot access device " + std::to_string(h.device));
      cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
      cudaGetLastError();
      base = (void*)(uintptr_t)h.ptr;
    } else {
      cudaError_t e = cudaIpcOpenMemHandle(&base, h.ipc, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) return fail(std::string("cudaIpcOpenMemHandle (rank ") + std::to_string(r) + "): " + cudaGetErrorString(e));
      m->xchg_mapped[r] = base;
    }
    X.data[r] = (double*)base;
    X.flags[r] = (unsigned long long*)((char*)base + xchg_data_doubles(nranks) * sizeof(double));
  }
  unsigned long long* tail = X.flags[rank] + (size_t)2 * nranks * XCHG_MAX_BLOCKS;
  X.seq = tail;
  X.ticket = (unsigned int*)(tail + 1);
  X.err = X.ticket + 1;
  m->xchg = X;
  m->xchg_on = true;
  return 0;
}

// ------------------------------------------------------------------------------------------------ MCMC driver
static int ensure_dep_sums(nimb200_model* m, int s) {  // block sums of the current state for sampler s's dependencies
  const nimb200_sampler& sp = m->samplers[s];
  for (int e = 0; e < sp.n_dep; e++) {
    int g = sp.dep_off + e, slot = m->dep_slot[g];
    if (m->dep_valid[slot]) continue;
    const nimb200_dep& dp = m->deps[g];
    HostSel hs = host_sel(m, dp, m->h_iarr);
    if (!hs.idx && hs.lo == 0 && hs.hi == m->decl[dp.decl].n && dp.decl < (int)m->lp_valid.size() && m->lp_valid[dp.decl]) {
      // initializeModel's calculate() of this declaration (P/R/initializeModel.R:78-96) was on this very state: reuse its sum
      CK(cudaMemcpyAsync(m->d_depcache + (size_t)slot * m->C, m->d_lpcache + (size_t)dp.decl * m->C, sizeof(double) * m->C,
                         cudaMemcpyDeviceToDevice, m->stream));
    } else if (launch_eval_sum(m, 0, dp.decl, hs.lo, hs.hi, hs.idx, m->d_depcache + (size_t)slot * m->C, nullptr, nullptr, -1, 0)) return 1;
    m->dep_valid[slot] = 1;
  }
  return 0;
}

static int run_sampler(nimb200_model* m, int s, uint64_t seed, uint64_t iter) {
  const nimb200_sampler& sp = m->samplers[s];
  int C = m->C;
  const int TB = 128;
  std::vector<int> refreshed;  // cache slots this sampler leaves valid although it touches their declarations
  if (sp.type == NIMB_SAMPLER_CATEGORICAL) {
    if (m->nranks > 1 && m->sweep_sharded[s]) return fail("categorical sampler over data-sharded dependents is not supported");
    size_t total = (size_t)sp.n_targets * C;
    k_categorical<<<(unsigned)((total + TB - 1) / TB), TB, 0, m->stream>>>(m->T, s, seed, iter);
    m->launches++;
  } else if (sp.type == NIMB_SAMPLER_SLICE) {
    if (m->nranks > 1 && m->sweep_sharded[s]) return fail("slice sampler over data-sharded dependents is not supported");
    size_t total = (size_t)sp.n_targets * C;
    k_slice<<<(unsigned)((total + TB - 1) / TB), TB, 0, m->stream>>>(m->T, s, seed, iter, m->sdev[s]);
    m->launches++;
  } else if (!m->wide[s] && m->bern_sweep[s] && !(m->nranks > 1 && m->sweep_sharded[s])) {
    size_t total = (size_t)sp.n_targets * C;
    const bool byp = m->bern_byp_slot[s][0] >= 0 && m->nranks <= 1;
    double* cz = byp ? m->d_bern_contrib[s] : nullptr;
    double* cy = byp ? m->d_bern_contrib[s] + total : nullptr;
    k_bernlin_binary_sweep<<<(unsigned)((total + BF_THREADS - 1) / BF_THREADS), BF_THREADS, 0, m->stream>>>(
        m->T, s, seed, iter, bern_args(m, m->deps[sp.dep_off].decl), bern_args(m, m->deps[sp.dep_off + 1].decl), cz, cy);
    m->launches++;
    if (byp) {
      const int chunks = (sp.n_targets + CS_ROWS - 1) / CS_ROWS;
      for (int k = 0; k < 2; k++) {
        k_colsum<<<dim3((C + 31) / 32, chunks), 256, 0, m->stream>>>(k == 0 ? cz : cy, sp.n_targets, C, m->d_partials);
        CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, m->d_partials, chunks, 1, C,
                                                              m->d_depcache + (size_t)m->bern_byp_slot[s][k] * C, nullptr));
        m->launches += 2;
        refreshed.push_back(m->bern_byp_slot[s][k]);
      }
    }
  } else if (!m->wide[s] && m->norm_sweep[s]) {
    // per-target (sum y, count) depend on the data only: computed on first use (data-sharded: this rank's slice, then one
    // all-reduce of 2 doubles per target) and kept; every later sweep is just the draws
    NormArgs na = norm_args(m, m->deps[sp.dep_off + 1].decl);
    if (!m->norm_gs_valid[s]) {
      k_norm_group_sums<<<sp.n_targets, 128, 0, m->stream>>>(m->T, s, na, m->d_norm_gs[s]);
      m->launches++;
      const bool shx = m->nranks > 1 && m->sweep_sharded[s] && comm_on(m);
      if (shx && allreduce_sum(m, m->d_norm_gs[s], (size_t)sp.n_targets * 2)) return 1;
      k_norm_group_ss<<<sp.n_targets, 128, 0, m->stream>>>(m->T, s, na, m->d_norm_gs[s], m->d_norm_gss[s]);
      m->launches++;
      if (shx && allreduce_sum(m, m->d_norm_gss[s], (size_t)sp.n_targets)) return 1;
      m->norm_gs_valid[s] = 1;
      // usable for whole-declaration block sums when every node belongs to a target and the statistics are finite (checked once,
      // outside any graph capture: the first sweep always runs through the ordinary launch path)
      m->norm_gss_ok[s] = 0;
      if (!getenv("NIMB200_NO_BYPRODUCT") && !(m->nranks > 1 && m->sweep_sharded[s] && !comm_on(m))) {
        std::vector<double> hg((size_t)sp.n_targets * 2), hs((size_t)sp.n_targets);
        CK(cudaMemcpyAsync(hg.data(), m->d_norm_gs[s], sizeof(double) * hg.size(), cudaMemcpyDeviceToHost, m->stream));
        CK(cudaMemcpyAsync(hs.data(), m->d_norm_gss[s], sizeof(double) * hs.size(), cudaMemcpyDeviceToHost, m->stream));
        CK(cudaStreamSynchronize(m->stream));
        double nodes = 0;
        bool fin = true;
        for (int t = 0; t < sp.n_targets; t++) { nodes += hg[2 * (size_t)t + 1]; fin = fin && std::isfinite(hg[2 * (size_t)t]) && std::isfinite(hs[t]); }
        m->norm_gss_ok[s] = fin && nodes == (double)m->decl[m->deps[sp.dep_off + 1].decl].n;
      }
    }
    size_t total = (size_t)sp.n_targets * C;
    CK(launch_pdl(k_norm_conj_draw, dim3((unsigned)((total + TB - 1) / TB)), dim3(TB), 0, m->stream, m->T, s, seed, iter, na, m->d_norm_gs[s]));
    m->launches++;
  } else if (!m->wide[s] && m->pois_sweep[s]) {
    size_t total = (size_t)sp.n_targets * C;
    // group-aligned shards: every rank updates the targets it owns, all of whose nodes are in its slice -- one phase, no exchange
    const bool sh = m->nranks > 1 && m->sweep_sharded[s] && !m->target_sharded[s];
    PoisArgs pa = pois_args(m, m->deps[sp.dep_off + 1].decl);
    if (!m->pois_ysum_valid[s]) {
      CK(cudaMemsetAsync(m->d_pois_ysum[s] + pois_stats_doubles((size_t)sp.n_targets) - 1, 0, sizeof(double), m->stream));   // max over targets
      k_pois_group_ysum<<<(sp.n_targets + 127) / 128, 128, 0, m->stream>>>(m->T, s, pa, (const PoisNode*)m->d_nodeconst[m->deps[sp.dep_off + 1].decl],
                                                                           m->d_pois_ysum[s]);
      m->launches++;
      if (sh && comm_on(m) && allreduce_sum(m, m->d_pois_ysum[s], (size_t)sp.n_targets * 3)) return 1;
      m->pois_ysum_valid[s] = 1;
    }
    const unsigned nb = (unsigned)((total + PF_THREADS - 1) / PF_THREADS);
    if (!sh) {
      // group-aligned shards keep the by-product: each rank adds up the shares of the targets it owns, one all-reduce of `chains`
      // doubles completes the cached block sum
      const bool own = m->nranks > 1 && m->target_sharded[s];
      const bool byp = m->pois_byp_slot[s] >= 0 && (m->nranks <= 1 || (own && comm_on(m)));
      if (pois_S_usable(m, s)) {   // rate sums from the group sums: exp(b0) S, no node is read
        if (!m->pois_S_valid[s]) {
          if (launch_pois_S(m, s, m->d_pois_S[s])) return 1;
          m->pois_S_valid[s] = 1;
        }
        const bool chain_level = m->pois_chain_level[s] != 0;   // the targets' prior parameters read nothing indexed by the node
        if (chain_level) {
          k_pois_chain_params<<<(C + 127) / 128, 128, 0, m->stream>>>(m->T, s, pa, m->d_pois_chainpar[s]);
          m->launches++;
        }
        k_pois_conj_sweep<4><<<nb, PF_THREADS, 0, m->stream>>>(m->T, s, seed, iter, pa, m->d_pois_ysum[s], m->d_pois_S[s],
                                                             byp ? m->d_pois_contrib[s] : nullptr,
                                                             chain_level ? (const double*)m->d_pois_chainpar[s] : (const double*)nullptr);
      } else
        k_pois_conj_sweep<0><<<nb, PF_THREADS, 0, m->stream>>>(m->T, s, seed, iter, pa, m->d_pois_ysum[s], nullptr,
                                                             byp ? m->d_pois_contrib[s] : nullptr);
      m->launches++;
      if (byp) {  // fixed-order sum over the targets -> the cached current-state block sum of the dependent declaration
        const int tdecl = m->deps[sp.dep_off].decl;
        const int t_lo = own ? m->shard_lo[tdecl] : 0, t_hi = own ? m->shard_hi[tdecl] : sp.n_targets;
        double* cache = m->d_depcache + (size_t)m->pois_byp_slot[s] * C;
        const int rows = std::max(0, t_hi - t_lo), cchunks = (rows + CS_ROWS - 1) / CS_ROWS;
        if (rows > 1024 && (size_t)cchunks * C <= m->partials_cap) {
          // many targets: coalesced column sums over chunks of rows first (k_reduce alone reads a column per warp, 32-byte
          // sectors for 8 bytes: 0.46 ms for the benchmark's 655 MB), then the fixed-order sum of the chunk partials
          k_colsum<<<dim3((C + 31) / 32, cchunks), 256, 0, m->stream>>>(m->d_pois_contrib[s] + (size_t)t_lo * C, rows, C, m->d_partials);
          CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, (const double*)m->d_partials, cchunks, 1, C, cache,
                        (double*)nullptr));
          m->launches += 2;
        } else {
          CK(launch_pdl(k_reduce, dim3((C * 32 + 127) / 128), dim3(128), 0, m->stream, (const double*)(m->d_pois_contrib[s] + (size_t)t_lo * C),
                        rows, 1, C, cache, (double*)nullptr));
          m->launches++;
        }
        if (own && allreduce_sum(m, cache, C)) return 1;
        refreshed.push_back(m->pois_byp_slot[s]);
      }
    } else {  // data-sharded: this rank's share of every (target, chain) rate contribution -> all-reduce -> draws
      if (total > m->sweepbuf_cap) return fail("internal: sweep buffer too small (poisson)");
      k_pois_conj_sweep<1><<<nb, PF_THREADS, 0, m->stream>>>(m->T, s, seed, iter, pa, m->d_pois_ysum[s], m->d_sweepbuf, nullptr);
      m->launches++;
      if (comm_on(m) && allreduce_sum(m, m->d_sweepbuf, total)) return 1;
      k_pois_conj_sweep<2><<<nb, PF_THREADS, 0, m->stream>>>(m->T, s, seed, iter, pa, m->d_pois_ysum[s], m->d_sweepbuf, nullptr);
      m->launches++;
    }
  } else if (!m->wide[s]) {
    size_t total = (size_t)sp.n_targets * C;
    // few (target, chain) pairs with many dependents each: a warp per pair
    bool warp_per_pair = total <= 65536 && m->sweep_deps[s] >= 8;
    if (warp_per_pair) total *= 32;
    unsigned nb = (unsigned)((total + TB - 1) / TB);
    if (m->nranks > 1 && m->sweep_sharded[s]) {  // partial sums -> all-reduce -> decide / draw
      size_t need = (size_t)sp.n_targets * C * 3;
      if (need > m->sweepbuf_cap) return fail("internal: sweep buffer too small");
      if (warp_per_pair) k_sweep<1, 32><<<nb, TB, 0, m->stream>>>(m->T, s, seed, iter, m->sdev[s], m->d_sweepbuf);
      else k_sweep<1, 1><<<nb, TB, 0, m->stream>>>(m->T, s, seed, iter, m->sdev[s], m->d_sweepbuf);
      m->launches++;
      if (comm_on(m) && allreduce_sum(m, m->d_sweepbuf, need)) return 1;
      k_sweep<2, 1><<<(unsigned)(((size_t)sp.n_targets * C + TB - 1) / TB), TB, 0, m->stream>>>(m->T, s, seed, iter, m->sdev[s], m->d_sweepbuf);
      m->launches++;
    } else {
      if (warp_per_pair) CK(launch_pdl(k_sweep<0, 32>, dim3(nb), dim3(TB), 0, m->stream, m->T, s, seed, iter, m->sdev[s], (double*)nullptr));
      else CK(launch_pdl(k_sweep<0, 1>, dim3(nb), dim3(TB), 0, m->stream, m->T, s, seed, iter, m->sdev[s], (double*)nullptr));
      m->launches++;
    }
  } else if (sp.type == NIMB_SAMPLER_CONJ_GAMMA || sp.type == NIMB_SAMPLER_CONJ_NORMAL || sp.type == NIMB_SAMPLER_CONJ_BETA) {
    const int gam = conj_kind(sp.type);
    for (int e = 0; e < sp.n_dep; e++) {
      const nimb200_dep& dp = m->deps[sp.dep_off + e];
      if (dp.stage != 1) continue;
      HostSel hs = host_sel(m, dp, m->h_iarr);
      if (launch_eval_sum(m, 2, dp.decl, hs.lo, hs.hi, hs.idx, m->d_newsum + ((size_t)e * 2) * C, m->d_newsum + ((size_t)e * 2 + 1) * C,
                          &dp, sp.target_slot, gam)) return 1;
    }
    CK(launch_pdl(k_conj_finish, dim3((C + TB - 1) / TB), dim3(TB), 0, m->stream, m->T, s, seed, iter, m->d_newsum));
    m->launches++;
  } else {
    int tries = sp.type == NIMB_SAMPLER_RW_BLOCK ? sp.tries : 1;
    int npair = (sp.d + 1) / 2;
    for (int tr = 0; tr < tries; tr++) {
      if (ensure_dep_sums(m, s)) return 1;
      // scalar RW: the target's own density under the proposal comes out of the proposal kernel (entry 0 = the target's node)
      const nimb200_dep& dp0 = m->deps[sp.dep_off];
      const HostSel hs0 = host_sel(m, dp0, m->h_iarr);
      const bool prior_fused = sp.type == NIMB_SAMPLER_RW && sp.n_targets == 1 && dp0.stage == 0 && !hs0.idx && hs0.hi - hs0.lo == 1 &&
                               m->decl[dp0.decl].value_kind == 1 && m->decl[dp0.decl].value_ref + hs0.lo == sp.target_slot &&
                               !(m->nranks > 1 && m->sharded[dp0.decl]) && !getenv("NIMB200_NO_FUSED_PRIOR");
      if (sp.type == NIMB_SAMPLER_RW)
        CK(launch_pdl(k_rw_propose, dim3((C + TB - 1) / TB), dim3(TB), 0, m->stream, m->T, s, seed, iter, m->sdev[s], m->d_tsave, m->d_pls,
                      prior_fused ? m->d_newsum : (double*)nullptr, hs0.lo));
      else
        k_block_propose<<<(C + 15) / 16, dim3(16, NIMB200_MAX_BLOCK_D), 0, m->stream>>>(m->T, s, seed, iter, tr, m->sdev[s], m->d_tsave);
      m->launches++;
      m->prop_lo = sp.target_slot; m->prop_hi = sp.target_slot + (sp.type == NIMB_SAMPLER_RW_BLOCK ? sp.d : 1);  // slots holding a proposal
      for (int e = prior_fused ? 1 : 0; e < sp.n_dep; e++) {
        const nimb200_dep& dp = m->deps[sp.dep_off + e];
        HostSel hs = host_sel(m, dp, m->h_iarr);
        if (launch_eval_sum(m, 0, dp.decl, hs.lo, hs.hi, hs.idx, m->d_newsum + (size_t)e * C, nullptr, nullptr, -1, 0)) return 1;
      }
      int call = sp.type == NIMB_SAMPLER_RW ? 1 : tr * (npair + 1) + npair;
      CK(launch_pdl(k_mh_finish, dim3((C + 31) / 32), dim3(32, MHF_J), 0, m->stream, m->T, s, seed, iter, call, m->sdev[s], m->d_newsum, m->d_depcache, m->d_dep_slot,
                                                           m->d_tsave, sp.type == NIMB_SAMPLER_RW ? m->d_pls : nullptr, m->d_accept));
      m->launches++;
      m->prop_lo = m->prop_hi = -1;
      for (size_t s2 = 0; s2 < m->samplers.size(); s2++) {  // a proposal that moved b1 of a Poisson family: accepted chains take the new group sums
        if (!m->pois_S_pending[s2]) continue;
        const int nT2 = m->samplers[s2].n_targets;
        k_pois_S_commit<<<dim3((C + 255) / 256, (nT2 + 63) / 64), 256, 0, m->stream>>>(m->d_pois_S[s2], m->d_pois_Snew[s2], m->d_accept, nT2, C);
        m->launches++;
        m->pois_S_pending[s2] = 0;
      }
      if (sp.type == NIMB_SAMPLER_RW_BLOCK && sp.adaptive) {
        m->ran[s]++;
        if (m->ran[s] % sp.adapt_interval == 0) {
          k_block_adapt<<<C, 128, 0, m->stream>>>(m->T, s, m->sdev[s]);
          m->launches++;
          m->ran[s] = 0;
        }
      }
    }
  }
  for (int g : m->invalidates[s]) m->dep_valid[g] = 0;
  for (int g : refreshed) m->dep_valid[g] = 1;
  {  // group sums follow b1 only through the wide MH path (proposal pass + commit); any other sampler moving b1 stales them
    const bool tracked = m->wide[s] && (sp.type == NIMB_SAMPLER_RW || sp.type == NIMB_SAMPLER_RW_BLOCK);
    const int lo_t = sp.target_slot, hi_t = sp.target_slot + (sp.type == NIMB_SAMPLER_RW_BLOCK ? sp.d : sp.n_targets);
    for (size_t s2 = 0; s2 < m->samplers.size() && !tracked; s2++) {
      if (!m->d_pois_S[s2]) continue;
      const int b1s = m->decl[m->deps[m->samplers[s2].dep_off + 1].decl].fam_arg[5];
      if (b1s >= lo_t && b1s < hi_t) m->pois_S_valid[s2] = 0;
    }
  }
  for (int e = 0; e < sp.n_dep && !m->lp_valid.empty(); e++) m->lp_valid[m->deps[sp.dep_off + e].decl] = 0;  // these sums have moved on
  {  // statistics of an i.i.d. gamma layer go stale when a sampler writes any of its values
    const int lo_t = sp.target_slot, hi_t = sp.target_slot + (sp.type == NIMB_SAMPLER_RW_BLOCK ? sp.d : sp.n_targets);
    for (size_t i = 0; i < m->gam_iid.size(); i++)
      if (m->gam_iid[i] && lo_t < m->decl[i].value_ref + m->decl[i].n && hi_t > m->decl[i].value_ref) m->gam_stats_valid[i] = 0;
  }
  CK(cudaGetLastError());
  return 0;
}

extern "C" int nimb200_mcmc_run(nimb200_model* m, int64_t niter, int64_t nburnin, int32_t thin, int32_t reset, uint64_t seed) {
  if (!m) return fail("null model");
  if (niter < 0) return fail("cannot specify niter < 0");
  if (nburnin < 0) return fail("cannot specify nburnin < 0");
  if (nburnin > niter) return fail("cannot specify nburnin > niter");
  if (thin < 1) return fail("cannot use thin < 1");
  if (!reset && nburnin != 0) return fail("cannot specify nburnin when using reset = FALSE.");
  CK(cudaSetDevice(m->device));
  int C = m->C, nmon = (int)m->monitors.size();
  if (reset) {
    m->rows = 0; m->iter_done = 0;
    if (m->waic) {
      m->waic_count = 0;
      for (size_t i = 0; i < m->decl.size(); i++)
        if (m->d_waic[i]) CK(cudaMemsetAsync(m->d_waic[i], 0, sizeof(double) * 4 * m->decl[i].n, m->stream));
    }
    for (size_t s = 0; s < m->samplers.size(); s++) {
      const nimb200_sampler& sp = m->samplers[s];
      int TT = sp.type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp.n_targets;
      size_t total = (size_t)TT * C;
      k_sampler_reset<<<(unsigned)((total + 127) / 128), 128, 0, m->stream>>>(m->T, (int)s, m->sdev[s], nullptr);
      if (sp.type == NIMB_SAMPLER_RW_BLOCK)
        k_block_reset<<<C, 64, 0, m->stream>>>(m->T, (int)s, m->sdev[s], sp.propcov_darr >= 0 ? m->d_darr[sp.propcov_darr] : nullptr);
      m->ran[s] = 0;
    }
    CK(cudaGetLastError());
  }
  // cached block sums survive a `reset = FALSE` continuation: nothing but the samplers has touched the chain state since the
  // last run (set_values / initialize_model / a new shard layout clear them themselves)
  if (reset) {
    std::fill(m->dep_valid.begin(), m->dep_valid.end(), 0);
    std::fill(m->pois_S_valid.begin(), m->pois_S_valid.end(), 0);
    std::fill(m->gam_stats_valid.begin(), m->gam_stats_valid.end(), 0);
  }
  m->prop_lo = m->prop_hi = -1;
  std::fill(m->pois_S_pending.begin(), m->pois_S_pending.end(), 0);
  int64_t addrows = (niter - nburnin) / thin;
  if (m->rows + addrows > m->cap) {  // resize(mvSamples, ...), MCMC_build.R:262
    int64_t ncap = m->rows + addrows;
    double* ns = nullptr;
    CK(cudaMalloc(&ns, std::max<size_t>(1, (size_t)ncap * nmon * C) * sizeof(double)));
    if (m->d_samples && m->rows)
      CK(cudaMemcpyAsync(ns, m->d_samples, sizeof(double) * (size_t)m->rows * nmon * C, cudaMemcpyDeviceToDevice, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    if (m->d_samples) cudaFree(m->d_samples);
    m->d_samples = ns; m->cap = ncap;
  }
  m->launches = 0;
  m->allreduces = 0;
  const int64_t rows_before = m->rows, iter0 = m->iter_done;
  {
    unsigned long long first = (unsigned long long)(m->iter_done + 1);
    CK(cudaMemcpyAsync(m->d_ctr, &first, sizeof first, cudaMemcpyHostToDevice, m->stream));
    RunParams rp{(long long)rows_before, (long long)iter0, (long long)nburnin, (int)thin, 0, m->d_samples};
    CK(cudaMemcpyAsync(m->d_runp, &rp, sizeof rp, cudaMemcpyHostToDevice, m->stream));
  }
  // One iteration = a fixed sequence of launches once the pattern of valid cached sums has settled, so it is captured into a
  // CUDA graph and replayed (launch latency is what bounds small models).  Iterations in which an RW_block adaptation kernel is
  // due run through the ordinary launch path; so does everything when a NCCL all-reduce is part of the sequence (the library's
  // own peer-memory all-reduce is an ordinary kernel and is captured).  What differs between calls (row offset, iteration
  // offset, burn-in, thinning, the sample buffer) is read from device memory, so the graph of an earlier call is replayed by
  // later ones whenever the seed and the cache pattern at the start of an iteration are those it was captured under.
  const char* genv = getenv("NIMB200_GRAPH");
  const bool want_graph = !(genv && genv[0] == '0') && !m->nccl_comm && niter >= 8 && !m->time_samplers;
  if (m->graph_exec && (!want_graph || m->waic || m->graph_seed != seed)) { cudaGraphExecDestroy(m->graph_exec); m->graph_exec = nullptr; }
  std::vector<char> valid_prev, dep_before, gam_before;
  // what decides the launch sequence of an iteration: which cached sums are valid, and which group sums are
  auto pattern = [&]() {
    std::vector<char> v = m->dep_valid;
    v.insert(v.end(), m->pois_S_valid.begin(), m->pois_S_valid.end());
    v.insert(v.end(), m->gam_stats_valid.begin(), m->gam_stats_valid.end());
    return v;
  };
  bool capture_failed = false;
  auto adapt_due = [&]() {
    for (size_t s = 0; s < m->samplers.size(); s++) {
      const nimb200_sampler& sp = m->samplers[s];
      if (sp.type != NIMB_SAMPLER_RW_BLOCK || !sp.adaptive) continue;
      for (int k = 1; k <= sp.tries; k++)
        if ((m->ran[s] + k) % sp.adapt_interval == 0) return true;
    }
    return false;
  };
  const long long waic_base = m->waic_count;
  auto one_iteration = [&](uint64_t git) -> int {
    for (size_t s = 0; s < m->samplers.size(); s++) {
      if (m->time_samplers) CK(cudaEventRecord(m->ev_s[2 * s], m->stream));
      if (run_sampler(m, (int)s, seed, git)) return 1;
      if (m->time_samplers) CK(cudaEventRecord(m->ev_s[2 * s + 1], m->stream));
    }
    if (m->time_samplers) {  // profiling mode: one synchronisation per iteration
      CK(cudaStreamSynchronize(m->stream));
      for (size_t s = 0; s < m->samplers.size(); s++) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, m->ev_s[2 * s], m->ev_s[2 * s + 1]));
        m->sampler_ms[s] += ms;
      }
    }
    if (m->waic) {  // before the monitor kernel, whose last block advances the iteration counter
      for (size_t i = 0; i < m->decl.size(); i++) {
        if (!m->d_waic[i] || m->decl[i].n == 0) continue;
        k_waic_update<<<(m->decl[i].n + 127) / 128, 128, 0, m->stream>>>(m->T, (int)i, m->d_waic[i], iter0, nburnin, waic_base);
        m->launches++;
      }
    }
    if (nmon > 0) {
      size_t total = (size_t)nmon * C;
      CK(launch_pdl(k_monitor_ctr, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, m->stream, m->T, (const int32_t*)m->d_mon, nmon,
                    (const RunParams*)m->d_runp, m->d_ctr, m->d_ticket));
      m->launches++;
    } else {
      k_bump_iter<<<1, 1, 0, m->stream>>>(m->d_ctr);
      m->launches++;
    }
    return 0;
  };
  CK(cudaEventRecord(m->ev0, m->stream));
  const bool small_path = m->small.n > 0 && m->nranks <= 1 && !m->waic && !m->time_samplers && nmon >= 0;
  if (small_path) {  // whole iterations inside one block, up to 256 of them per launch
    int threads = 32;
    for (int j = 0; j < m->small.n; j++) {
      const size_t tot = (size_t)m->samplers[m->small.sid[j]].n_targets * C * m->small.lanes[j];
      threads = std::max(threads, (int)std::min<size_t>(1024, (tot + 31) / 32 * 32));
    }
    for (int64_t done = 0; done < niter;) {
      const int n_it = (int)std::min<int64_t>(256, niter - done);
      k_small_iters<<<1, threads, 0, m->stream>>>(m->T, m->small, m->d_sdev, seed, n_it, (const int32_t*)m->d_mon, nmon,
                                                  (const long long*)m->d_runp, m->d_samples, m->d_ctr);
      m->launches++;
      done += n_it;
    }
    CK(cudaGetLastError());
  }
  for (int64_t it = 1; it <= niter && !small_path; it++) {
    uint64_t git = (uint64_t)(m->iter_done + it);
    bool due = adapt_due();
    if (m->graph_exec && !due && pattern() == m->graph_pattern) {
      CK(cudaGraphLaunch(m->graph_exec, m->stream));
      m->launches += m->graph_launches;
      m->allreduces += m->graph_allreduces;
      for (size_t s = 0; s < m->samplers.size(); s++)
        if (m->samplers[s].type == NIMB_SAMPLER_RW_BLOCK && m->samplers[s].adaptive) m->ran[s] += m->samplers[s].tries;
      continue;
    }
    // capture once the pattern repeats from one iteration to the next (and an older graph, if any, does not fit it)
    bool capture = want_graph && !capture_failed && !due && it >= 3 && valid_prev == pattern();
    valid_prev = pattern();
    dep_before = m->dep_valid;
    gam_before = m->gam_stats_valid;
    if (capture) {
      std::vector<int> ran_before = m->ran;
      int64_t l0 = m->launches, a0 = m->allreduces;
      cudaGraph_t g = nullptr;
      if (m->graph_exec) { cudaGraphExecDestroy(m->graph_exec); m->graph_exec = nullptr; }
      bool ok = cudaStreamBeginCapture(m->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
      if (ok && one_iteration(git)) ok = false;
      if (cudaStreamEndCapture(m->stream, &g) != cudaSuccess || !g) ok = false;
      if (ok && pattern() != valid_prev) ok = false;  // the cache pattern must repeat for the replay to be valid
      if (ok && cudaGraphInstantiate(&m->graph_exec, g, 0) != cudaSuccess) { m->graph_exec = nullptr; ok = false; }
      if (g) cudaGraphDestroy(g);
      m->graph_launches = m->launches - l0;
      m->graph_allreduces = m->allreduces - a0;
      m->launches = l0;
      m->allreduces = a0;
      if (ok) {
        m->graph_pattern = valid_prev;
        m->graph_seed = seed;
        CK(cudaGraphLaunch(m->graph_exec, m->stream));  // capturing did not execute anything
        m->launches += m->graph_launches;
        m->allreduces += m->graph_allreduces;
        continue;
      }
      cudaGetLastError();
      m->ran = ran_before;  // nothing ran: fall back to ordinary launches for good
      m->dep_valid = dep_before;
      m->gam_stats_valid = gam_before;
      if (m->graph_exec) { cudaGraphExecDestroy(m->graph_exec); m->graph_exec = nullptr; }
      capture_failed = true;
    }
    if (one_iteration(git)) return 1;
  }
  m->rows += (niter - nburnin) / thin;
  CK(cudaEventRecord(m->ev1, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  CK(cudaGetLastError());
  if (m->xchg_on) {
    unsigned int bad = 0;
    CK(cudaMemcpy(&bad, m->xchg.err, sizeof bad, cudaMemcpyDeviceToHost));
    if (bad) return fail("peer all-reduce timed out: a rank of the data-sharded group did not arrive (results are NaN)");
  }
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, m->ev0, m->ev1));
  m->last_ms = ms; m->last_launches = m->launches;
  m->iter_done += niter;
  if (m->waic) m->waic_count += (int64_t)(niter - nburnin) * C;
  return 0;
}

extern "C" int nimb200_mcmc_samples(nimb200_model* m, double* out, int64_t* nrows_out) {
  if (!m) return fail("null model");
  if (nrows_out) *nrows_out = m->rows;
  if (!out) return 0;
  CK(cudaSetDevice(m->device));
  int nmon = (int)m->monitors.size();
  size_t total = (size_t)m->rows * nmon * m->C;
  if (total == 0) return 0;
  double* tmp = nullptr;
  CK(cudaMalloc(&tmp, sizeof(double) * total));
  k_samples_out<<<(unsigned)((total + 255) / 256), 256, 0, m->stream>>>(m->d_samples, tmp, m->rows, nmon, m->C);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, tmp, sizeof(double) * total, cudaMemcpyDeviceToHost, m->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
  cudaFree(tmp);
  if (e != cudaSuccess) return fail(std::string("CUDA error ") + cudaGetErrorString(e));
  return 0;
}

// ------------------------------------------------------------------------------------------------ samplesSummary on the device
// samplesSummary (P/R/MCMC_utils.R:315-341): Mean, Median, St.Dev., 2.5 % and 97.5 % quantiles of every monitored column, over
// the rows of the chains [chain_lo, chain_hi) pooled -- without bringing mvSamples to the host.  Exact order statistics by
// radix selection on order-preserving 64-bit keys (six passes of 12 bits, one shared-memory histogram per column block and
// wanted rank), quantiles by R's default (type 7) interpolation; moments by fixed-order two-stage sums.
static const int SUM_RANKS = 3;                 // ranks floor((n-1) p) for p = 0.025, 0.5, 0.975
static const int SUM_BINS = 4096;
__device__ __forceinline__ unsigned long long ord_key(double x) {
  unsigned long long u = (unsigned long long)__double_as_longlong(x);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__host__ __device__ inline double ord_unkey(unsigned long long k) {
  unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  double d;
  memcpy(&d, &u, sizeof d);
  return d;
}
// element e of a column's pooled sample: row e / nc, chain c_lo + e % nc;  samples[row][col][chain]
struct SumView { const double* s; long long rows; int nmon, C, c_lo, nc; };
__device__ __forceinline__ double sum_elem(const SumView& v, int col, long long e) {
  const long long row = e / v.nc;
  const int c = v.c_lo + (int)(e % v.nc);
  return v.s[((size_t)row * v.nmon + col) * v.C + c];
}
// grid (chunks, columns): pass = 0: sum; pass = 1: sum of squared deviations from mean[col].  partial[chunk][col]
__global__ void k_sum_moments(SumView v, int pass, const double* __restrict__ mean, double* __restrict__ partial) {
  __shared__ double red[8];
  const int col = blockIdx.y;
  const long long n = v.rows * v.nc, per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = blockIdx.x * per, hi = min(n, lo + per);
  const double mu = pass ? mean[col] : 0.0;
  double acc = 0.0;
  for (long long e = lo + threadIdx.x; e < hi; e += blockDim.x) {
    const double x = sum_elem(v, col, e) - mu;
    acc += pass ? x * x : x;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += red[w];
    partial[(size_t)blockIdx.x * gridDim.y + col] = t;
  }
}
// one 12-bit digit of the radix selection: for each wanted rank r of the column, count the elements whose key agrees with
// prefix[col][r] above bit (shift + 12) by their digit at `shift`.  hist[col][r][bin]
__global__ void k_sum_hist(SumView v, int shift, int cmp_shift, unsigned int mask, int first, const unsigned long long* __restrict__ prefix,
                           unsigned int* __restrict__ hist) {
  extern __shared__ unsigned int sum_sh[];   // [SUM_RANKS][SUM_BINS]
  unsigned int* sh = sum_sh;
  for (int i = threadIdx.x; i < SUM_RANKS * SUM_BINS; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int col = blockIdx.y;
  const long long n = v.rows * v.nc, per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = blockIdx.x * per, hi = min(n, lo + per);
  unsigned long long pf[SUM_RANKS];
#pragma unroll
  for (int r = 0; r < SUM_RANKS; r++) pf[r] = prefix[(size_t)col * SUM_RANKS + r];
  for (long long e = lo + threadIdx.x; e < hi; e += blockDim.x) {
    const unsigned long long k = ord_key(sum_elem(v, col, e));
    const unsigned int bin = (unsigned int)(k >> shift) & mask;
#pragma unroll
    for (int r = 0; r < SUM_RANKS; r++)
      if (first || (k >> cmp_shift) == (pf[r] >> cmp_shift)) atomicAdd(&sh[r * SUM_BINS + bin], 1u);
  }
  __syncthreads();
  unsigned int* g = hist + (size_t)col * SUM_RANKS * SUM_BINS;
  for (int i = threadIdx.x; i < SUM_RANKS * SUM_BINS; i += blockDim.x)
    if (sh[i]) atomicAdd(&g[i], sh[i]);
}
// for each (column, rank): how many elements are <= the selected value, and the smallest key above it (the next order statistic)
__global__ void k_sum_next(SumView v, const unsigned long long* __restrict__ sel, unsigned long long* __restrict__ count_le,
                           unsigned long long* __restrict__ min_gt) {
  const int col = blockIdx.y;
  const long long n = v.rows * v.nc, per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = blockIdx.x * per, hi = min(n, lo + per);
  unsigned long long sk[SUM_RANKS], cnt[SUM_RANKS], mg[SUM_RANKS];
#pragma unroll
  for (int r = 0; r < SUM_RANKS; r++) { sk[r] = sel[(size_t)col * SUM_RANKS + r]; cnt[r] = 0; mg[r] = ~0ull; }
  for (long long e = lo + threadIdx.x; e < hi; e += blockDim.x) {
    const unsigned long long k = ord_key(sum_elem(v, col, e));
#pragma unroll
    for (int r = 0; r < SUM_RANKS; r++) { if (k <= sk[r]) cnt[r]++; else mg[r] = min(mg[r], k); }
  }
#pragma unroll
  for (int r = 0; r < SUM_RANKS; r++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      cnt[r] += __shfl_xor_sync(0xffffffffu, cnt[r], o);
      mg[r] = min(mg[r], __shfl_xor_sync(0xffffffffu, mg[r], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicAdd(&count_le[(size_t)col * SUM_RANKS + r], cnt[r]);
      atomicMin(&min_gt[(size_t)col * SUM_RANKS + r], mg[r]);
    }
  }
}

// MCMCsaveHistory (nimbleOptions(MCMCsaveHistory = TRUE); P/R/MCMC_samplers.R:394-399, 726-739, getScaleHistory :421-448): once
// enabled, every adaptation of every RW / RW_block sampler records the scale it used, the interval's acceptance rate and (RW_block,
// d <= 10 = maxDimCovHistory) the proposal covariance it used, for up to max_adaptations adaptations since the last reset.
extern "C" int nimb200_sampler_history_enable(nimb200_model* m, int32_t max_adaptations) {
  if (!m || max_adaptations < 1) return fail("bad argument");
  CK(cudaSetDevice(m->device));
  for (size_t s = 0; s < m->samplers.size(); s++) {
    const nimb200_sampler& sp = m->samplers[s];
    if (sp.type != NIMB_SAMPLER_RW && sp.type != NIMB_SAMPLER_RW_BLOCK) continue;
    SamplerDev& S = m->sdev[s];
    const int TT = sp.type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp.n_targets;
    const int K = 2 + (sp.type == NIMB_SAMPLER_RW_BLOCK && sp.d <= 10 ? sp.d * sp.d : 0);
    const size_t n = (size_t)max_adaptations * K * TT * m->C;
    if (dev_alloc(m, &S.hist, n)) return 1;
    CK(cudaMemsetAsync(S.hist, 0, sizeof(double) * n, m->stream));
    S.hist_cap = max_adaptations; S.hist_K = K;
  }
  if (m->d_sdev) CK(cudaMemcpyAsync(m->d_sdev, m->sdev.data(), sizeof(SamplerDev) * m->sdev.size(), cudaMemcpyHostToDevice, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  if (m->graph_exec) { cudaGraphExecDestroy(m->graph_exec); m->graph_exec = nullptr; }   // the sampler descriptors are kernel arguments
  return 0;
}
// what: 0 scaleHistory, 1 acceptanceHistory -> out[adaptation][T][C]; 2 propCovHistory -> out[adaptation][d*d][C] (column-major
// d x d per adaptation and chain).  *n_adapt_out = adaptations recorded since the last reset; out may be NULL to query it.
extern "C" int nimb200_sampler_history(nimb200_model* m, int32_t s, int32_t what, double* out, int64_t cap, int32_t* n_adapt_out) {
  if (!m) return fail("null model");
  if (s < 0 || s >= (int)m->samplers.size()) return fail("bad sampler index");
  CK(cudaSetDevice(m->device));
  const nimb200_sampler& sp = m->samplers[s];
  const SamplerDev& S = m->sdev[s];
  if (!S.hist) return fail("Please enable the sampler histories (MCMCsaveHistory) before running the MCMC.");
  const int TT = sp.type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp.n_targets;
  int adapted = 0;
  CK(cudaMemcpyAsync(&adapted, S.timesAdapted, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  adapted = std::min(adapted, S.hist_cap);
  if (n_adapt_out) *n_adapt_out = adapted;
  if (!out) return 0;
  if (what < 0 || what > 2 || (what == 2 && S.hist_K <= 2)) return fail("no such history for this sampler (propCov histories are kept for RW_block with d <= 10)");
  const size_t TC = (size_t)TT * m->C, per = what == 2 ? (size_t)(S.hist_K - 2) * TC : TC;
  if ((int64_t)(per * adapted) > cap) return fail("output too small");
  for (int a = 0; a < adapted; a++)
    CK(cudaMemcpyAsync(out + (size_t)a * per, S.hist + ((size_t)a * S.hist_K + (what == 2 ? 2 : what)) * TC, sizeof(double) * per,
                       cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  return 0;
}

extern "C" int nimb200_samples_summary(nimb200_model* m, int32_t chain_lo, int32_t chain_hi, double* out) {
  if (!m || !out) return fail("null argument");
  if (chain_lo < 0 || chain_hi > m->C || chain_lo >= chain_hi) return fail("bad chain range");
  CK(cudaSetDevice(m->device));
  const int nmon = (int)m->monitors.size();
  const long long n = (long long)m->rows * (chain_hi - chain_lo);
  if (nmon == 0) return 0;
  if (n < 1) return fail("samplesSummary: no samples");
  SumView v{m->d_samples, (long long)m->rows, nmon, m->C, chain_lo, chain_hi - chain_lo};
  const int chunks = (int)std::max<long long>(1, std::min<long long>(64, n / 4096 + 1));
  const size_t NR = (size_t)nmon * SUM_RANKS;
  double *d_part = nullptr, *d_mean = nullptr;
  unsigned long long *d_prefix = nullptr, *d_cnt = nullptr, *d_min = nullptr;
  unsigned int* d_hist = nullptr;
  cudaError_t e = cudaSuccess;
  auto ck = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
  ck(cudaMalloc(&d_part, sizeof(double) * (size_t)chunks * nmon));
  ck(cudaMalloc(&d_mean, sizeof(double) * nmon));
  ck(cudaMalloc(&d_prefix, sizeof(unsigned long long) * NR));
  ck(cudaMalloc(&d_cnt, sizeof(unsigned long long) * NR));
  ck(cudaMalloc(&d_min, sizeof(unsigned long long) * NR));
  ck(cudaMalloc(&d_hist, sizeof(unsigned int) * NR * SUM_BINS));
  std::vector<double> part((size_t)chunks * nmon), mean(nmon), sd(nmon);
  std::vector<unsigned long long> prefix(NR, 0), want(NR), cnt(NR), mn(NR);
  std::vector<unsigned int> hist(NR * SUM_BINS);
  const double ps[SUM_RANKS] = {0.025, 0.5, 0.975};
  if (e == cudaSuccess) {
    // moments: fixed-order partial sums, finished on the host in chunk order
    for (int pass = 0; pass < 2 && e == cudaSuccess; pass++) {
      k_sum_moments<<<dim3(chunks, nmon), 256, 0, m->stream>>>(v, pass, d_mean, d_part);
      ck(cudaMemcpyAsync(part.data(), d_part, sizeof(double) * part.size(), cudaMemcpyDeviceToHost, m->stream));
      ck(cudaStreamSynchronize(m->stream));
      for (int j = 0; j < nmon; j++) {
        double t = 0.0;
        for (int b = 0; b < chunks; b++) t += part[(size_t)b * nmon + j];
        if (pass == 0) mean[j] = t / (double)n; else sd[j] = n > 1 ? std::sqrt(t / (double)(n - 1)) : NAN;
      }
      if (pass == 0) ck(cudaMemcpyAsync(d_mean, mean.data(), sizeof(double) * nmon, cudaMemcpyHostToDevice, m->stream));
    }
    // order statistics floor((n-1) p): radix selection, most significant digit first (digits at bits 52, 40, 28, 16, 4; then the low 4 bits)
    for (size_t q = 0; q < NR; q++) want[q] = (unsigned long long)std::floor((double)(n - 1) * ps[q % SUM_RANKS]);
    const int shifts[6] = {52, 40, 28, 16, 4, 0};   // five 12-bit digits, then the low 4 bits
    for (int ps_ = 0; ps_ < 6 && e == cudaSuccess; ps_++) {
      const int shift = shifts[ps_];
      const bool last = ps_ == 5;
      ck(cudaMemsetAsync(d_hist, 0, sizeof(unsigned int) * hist.size(), m->stream));
      ck(cudaMemcpyAsync(d_prefix, prefix.data(), sizeof(unsigned long long) * NR, cudaMemcpyHostToDevice, m->stream));
      k_sum_hist<<<dim3(chunks, nmon), 256, sizeof(unsigned int) * SUM_RANKS * SUM_BINS, m->stream>>>(
          v, shift, last ? 4 : (ps_ == 0 ? 63 : shift + 12), last ? 15u : (unsigned)(SUM_BINS - 1), ps_ == 0, d_prefix, d_hist);
      ck(cudaGetLastError());
      ck(cudaMemcpyAsync(hist.data(), d_hist, sizeof(unsigned int) * hist.size(), cudaMemcpyDeviceToHost, m->stream));
      ck(cudaStreamSynchronize(m->stream));
      for (size_t q = 0; q < NR && e == cudaSuccess; q++) {
        const unsigned int* h = hist.data() + q * SUM_BINS;
        unsigned long long acc = 0;
        int bin = 0;
        for (; bin < SUM_BINS; bin++) { if (acc + h[bin] > want[q]) break; acc += h[bin]; }
        if (bin == SUM_BINS) bin = SUM_BINS - 1;
        want[q] -= acc;
        prefix[q] |= (unsigned long long)bin << shift;
      }
    }
    // the next order statistic above each selected value (type-7 interpolation needs ranks lo and lo + 1)
    std::fill(cnt.begin(), cnt.end(), 0ull);
    std::fill(mn.begin(), mn.end(), ~0ull);
    ck(cudaMemcpyAsync(d_prefix, prefix.data(), sizeof(unsigned long long) * NR, cudaMemcpyHostToDevice, m->stream));
    ck(cudaMemcpyAsync(d_cnt, cnt.data(), sizeof(unsigned long long) * NR, cudaMemcpyHostToDevice, m->stream));
    ck(cudaMemcpyAsync(d_min, mn.data(), sizeof(unsigned long long) * NR, cudaMemcpyHostToDevice, m->stream));
    k_sum_next<<<dim3(chunks, nmon), 256, 0, m->stream>>>(v, d_prefix, d_cnt, d_min);
    ck(cudaGetLastError());
    ck(cudaMemcpyAsync(cnt.data(), d_cnt, sizeof(unsigned long long) * NR, cudaMemcpyDeviceToHost, m->stream));
    ck(cudaMemcpyAsync(mn.data(), d_min, sizeof(unsigned long long) * NR, cudaMemcpyDeviceToHost, m->stream));
    ck(cudaStreamSynchronize(m->stream));
  }
  cudaFree(d_part); cudaFree(d_mean); cudaFree(d_prefix); cudaFree(d_cnt); cudaFree(d_min); cudaFree(d_hist);
  if (e != cudaSuccess) return fail(std::string("CUDA error ") + cudaGetErrorString(e));
  for (int j = 0; j < nmon; j++) {
    double qv[SUM_RANKS];
    for (int r = 0; r < SUM_RANKS; r++) {
      const size_t q = (size_t)j * SUM_RANKS + r;
      const double h = (double)(n - 1) * ps[r];
      const long long lo = (long long)std::floor(h);
      const double xlo = ord_unkey(prefix[q]);
      // rank lo + 1: the same value when duplicates reach past rank lo, else the smallest larger element
      const double xhi = (cnt[q] > (unsigned long long)(lo + 1) || mn[q] == ~0ull) ? xlo : ord_unkey(mn[q]);
      qv[r] = xlo + (h - (double)lo) * (xhi - xlo);
    }
    out[0 * (size_t)nmon + j] = mean[j];
    out[1 * (size_t)nmon + j] = qv[1];
    out[2 * (size_t)nmon + j] = sd[j];
    out[3 * (size_t)nmon + j] = qv[0];
    out[4 * (size_t)nmon + j] = qv[2];
  }
  return 0;
}

extern "C" int nimb200_mcmc_last_allreduces(nimb200_model* m, int64_t* count_out, int32_t* graph_replayed_out) {
  if (!m) return fail("null model");
  if (count_out) *count_out = m->allreduces;
  if (graph_replayed_out) *graph_replayed_out = m->graph_exec ? 1 : 0;
  return 0;
}

extern "C" int nimb200_mcmc_last_timing(nimb200_model* m, double* ms_out, int64_t* launches_out) {
  if (!m) return fail("null model");
  if (ms_out) *ms_out = m->last_ms;
  if (launches_out) *launches_out = m->last_launches;
  return 0;
}

__global__ void k_i2d(const int32_t* src, double* dst, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = (double)src[i];
}

extern "C" int nimb200_sampler_state(nimb200_model* m, int32_t s, int32_t what, double* out, int64_t cap) {
  if (!m || !out) return fail("null argument");
  if (s < 0 || s >= (int)m->samplers.size()) return fail("bad sampler index");
  CK(cudaSetDevice(m->device));
  const nimb200_sampler& sp = m->samplers[s];
  int TT = sp.type == NIMB_SAMPLER_RW_BLOCK ? 1 : sp.n_targets;
  size_t n;
  const double* src = nullptr;
  double* tmp = nullptr;
  if (what == 0) { n = (size_t)TT * m->C; src = m->sdev[s].scale; }
  else if (what == 1) {
    n = (size_t)TT * m->C;
    CK(cudaMalloc(&tmp, sizeof(double) * n));
    k_i2d<<<(unsigned)((n + 255) / 256), 256, 0, m->stream>>>(m->sdev[s].timesAdapted, tmp, n);
    src = tmp;
  } else if (what == 2 && sp.type == NIMB_SAMPLER_RW_BLOCK) { n = (size_t)sp.d * sp.d * m->C; src = m->sdev[s].propCov; }
  else return fail("bad state selector");
  if ((int64_t)n > cap) { if (tmp) cudaFree(tmp); return fail("output too small"); }
  cudaError_t e = cudaMemcpyAsync(out, src, sizeof(double) * n, cudaMemcpyDeviceToHost, m->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
  if (tmp) cudaFree(tmp);
  if (e != cudaSuccess) return fail(std::string("CUDA error ") + cudaGetErrorString(e));
  return 0;
}

extern "C" int nimb200_mcmc_set_timing(nimb200_model* m, int32_t enable) {
  if (!m) return fail("null model");
  CK(cudaSetDevice(m->device));
  if (enable && m->ev_s.empty()) {
    m->ev_s.resize(2 * m->samplers.size());
    for (auto& e : m->ev_s) CK(cudaEventCreate(&e));
  }
  if (enable) m->sampler_ms.assign(m->samplers.size(), 0.0);
  m->time_samplers = enable != 0;
  return 0;
}

extern "C" int nimb200_mcmc_sampler_times(nimb200_model* m, double* out_ms, int32_t n) {
  if (!m || !out_ms) return fail("null argument");
  if (n < (int)m->samplers.size()) return fail("output too small");
  for (size_t s = 0; s < m->samplers.size(); s++) out_ms[s] = s < m->sampler_ms.size() ? m->sampler_ms[s] : 0.0;
  return 0;
}

extern "C" int nimb200_waic_enable(nimb200_model* m, int32_t enable) {
  if (!m) return fail("null model");
  CK(cudaSetDevice(m->device));
  if (enable && m->d_waic.empty()) {
    m->d_waic.assign(m->decl.size(), nullptr);
    for (size_t i = 0; i < m->decl.size(); i++) {
      if (m->decl[i].value_kind != 0 || m->decl[i].n == 0) continue;
      if (dev_alloc(m, &m->d_waic[i], (size_t)4 * m->decl[i].n)) return 1;
      CK(cudaMemsetAsync(m->d_waic[i], 0, sizeof(double) * 4 * m->decl[i].n, m->stream));
    }
    m->waic_count = 0;
  }
  m->waic = enable != 0;
  return 0;
}

// the sums behind WAIC over the data nodes this rank accounts for: its slice of a sharded declaration; a replicated one on rank 0
static int waic_sums(nimb200_model* m, double* logsum_out, double* ssq_out, long* nnodes_out) {
  if (m->d_waic.empty()) return fail("online WAIC was not enabled (nimb200_waic_enable)");
  CK(cudaSetDevice(m->device));
  CK(cudaStreamSynchronize(m->stream));
  double logsum = 0.0, ssq = 0.0;
  long nnodes = 0;
  std::vector<double> h;
  for (size_t i = 0; i < m->decl.size(); i++) {
    if (!m->d_waic[i]) continue;
    const size_t n = (size_t)m->decl[i].n;
    size_t lo = 0, hi = n;
    if (m->nranks > 1) {
      if (m->sharded[i]) { lo = (size_t)m->shard_lo[i]; hi = (size_t)m->shard_hi[i]; }
      else if (m->rank != 0) hi = 0;
    }
    h.resize(4 * n);
    CK(cudaMemcpy(h.data(), m->d_waic[i], sizeof(double) * 4 * n, cudaMemcpyDeviceToHost));
    for (size_t j = lo; j < hi; j++) { logsum += h[j] + log(h[n + j]); ssq += h[3 * n + j]; }  // MCMC_WAIC.R:276-278
    nnodes += (long)(hi > lo ? hi - lo : 0);
  }
  *logsum_out = logsum; *ssq_out = ssq; *nnodes_out = nnodes;
  return 0;
}

extern "C" int nimb200_waic_partial(nimb200_model* m, double* out4) {
  if (!m || !out4) return fail("null argument");
  double logsum, ssq;
  long nnodes;
  if (waic_sums(m, &logsum, &ssq, &nnodes)) return 1;
  out4[0] = logsum; out4[1] = ssq; out4[2] = (double)nnodes; out4[3] = (double)m->waic_count;
  return 0;
}

extern "C" int nimb200_waic_get(nimb200_model* m, double* out4) {
  if (!m || !out4) return fail("null argument");
  if (m->nranks > 1) return fail("data-sharded model: sum nimb200_waic_partial over the ranks of the group (nimble_b200.parallel.waic_sharded)");
  const double cnt = (double)m->waic_count;
  double logsum, ssq;
  long nnodes;
  if (waic_sums(m, &logsum, &ssq, &nnodes)) return 1;
  if (m->waic_count < 2) { out4[0] = out4[1] = out4[2] = NAN; out4[3] = cnt; return 0; }
  const double lppd = logsum - (double)nnodes * log(cnt), pwaic = ssq / (cnt - 1.0);
  out4[0] = -2.0 * (lppd - pwaic); out4[1] = lppd; out4[2] = pwaic; out4[3] = cnt;
  return 0;
}

extern "C" int nimb200_initialize_model(nimb200_model* m, const int32_t* decl_order, int32_t n_order, uint64_t seed) {
  if (!m || (!decl_order && n_order > 0) || n_order < 0) return fail("bad argument");
  CK(cudaSetDevice(m->device));
  int* d_flag = (int*)m->d_ticket;  // free between runs: the monitor kernel leaves it at zero
  CK(cudaMemsetAsync(d_flag, 0, sizeof(int), m->stream));
  for (int q = 0; q < n_order; q++) {
    int di = decl_order[q];
    if (di < 0 || di >= (int)m->decl.size()) return fail("declaration index out of range");
    const nimb200_decl& d = m->decl[di];
    if (d.value_kind != 1 || d.n == 0) continue;
    size_t total = (size_t)d.n * m->C;
    k_init_simulate<<<(unsigned)((total + 127) / 128), 128, 0, m->stream>>>(m->T, di, seed, d_flag);
  }
  int flag = 0;
  CK(cudaMemcpyAsync(&flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  CK(cudaMemsetAsync(d_flag, 0, sizeof(int), m->stream));
  if (flag) return fail("initializeModel: a missing initial value cannot be simulated (declaration " + std::to_string(flag - 1) +
                        ": only dnorm, dgamma, dbeta, dexp, dlogis, dbin and dpois (mean <= 500) nodes can)");
  std::fill(m->dep_valid.begin(), m->dep_valid.end(), 0);
  std::fill(m->lp_valid.begin(), m->lp_valid.end(), 0);
  std::fill(m->gam_stats_valid.begin(), m->gam_stats_valid.end(), 0);
  std::fill(m->pois_S_valid.begin(), m->pois_S_valid.end(), 0);
  return 0;
}

extern "C" int nimb200_draws(uint64_t seed, int32_t chain, int64_t iter, int32_t sampler, int32_t target, int32_t kind,
                             double a, int32_t n, double* out) {
  if (!out || n < 0) return fail("bad argument");
  if (n == 0) return 0;
  double* d = nullptr;
  CK(cudaMalloc(&d, sizeof(double) * n));
  k_draws<<<(n + 127) / 128, 128>>>(seed, chain, (uint64_t)iter, sampler, target, kind, a, n, d);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpy(out, d, sizeof(double) * n, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return fail(std::string("CUDA error ") + cudaGetErrorString(e));
  return 0;
}

// fused_pois.cuh -- fused kernels for  y[i] ~ dpois(lam[G(i)] * exp(b0 + b1 * x[i]))  (BASELINE config 5: Poisson GLMM).
// Included by engine.cu after the interpreter helpers (sel_of, get_param, make_key, rgamma1).
//
// The generic interpreter spends > 1000 instructions per evaluation on this declaration: the postfix program, CUDA's
// exp, and nmath's dpois_raw with its stirlerr table look-up (divergent constant-memory index), log(2 pi y) and the
// bd0 Taylor loop -- most of which depends on the data only.  With mu = lam exp(eta), eta = b0 + b1 x:
//   dpois_raw(y, mu) = -stirlerr(y) - 0.5 log(2 pi y) - (y log(y / mu) + mu - y)
//                    = A[i] + y (log lam + eta) - mu,      A[i] = -stirlerr(y) - 0.5 log(2 pi y) + y - y log y   (data only)
// so per (node, chain) only one table-driven exp (fastmath.cuh) remains: no logarithm, no reciprocal, no series, no
// divergence; log lam is taken when the group changes, and y = 0 needs no special case (A = 0, lp = -mu).
// nmath evaluates bd0 by a Taylor series when y ~ mu to keep its RELATIVE accuracy; this form has absolute error
// ~ 2e-16 y log y, which relative to lp (|lp| >= 0.5 log(2 pi y) for y >= 1) stays below 4e-13 for counts up to 1024:
// inside the 1e-12 parity bar (tests/test_gpu_parity.py::test_pois_family_*).  Anything else (y > 1024, lam outside
// [1e-150, 1e150], NaN, |eta| >= 300) has the whole run of nodes redone, the exact generic path for those elements.
// Round 2: an MCMC iteration no longer walks the nodes at all while proposals are of posterior width -- block sums and the
// conjugate sweep come from per-group sums S_g(b1) = sum exp(b1 x), themselves evaluated from 25 data-only moment coefficients
// per group (k_pois_S_series); see the comments at k_pois_conj_sweep and k_pois_S_series.
#pragma once

namespace nimb {

struct PoisArgs {
  const double* y;
  const double* x;
  const int32_t* gidx;   // index array (mode NIMB_IDX_IARR) or null
  int mode, arg;         // index map i -> lam element
  int lam_slot, b0_slot, b1_slot;
  int n;
};

__device__ __forceinline__ int pois_lam_index(const PoisArgs& a, int i) {
  switch (a.mode) {
    case NIMB_IDX_I: return i;
    case NIMB_IDX_DIV: return i / a.arg;
    case NIMB_IDX_MOD: return i % a.arg;
    default: return a.gidx[i];
  }
}

// one staged node: 32 bytes, built once per model by k_pois_pack (x is NaN for nodes that must take the exact path)
struct __align__(16) PoisNode {
  double y, x, A;
  int g, pad;
};
static const double PF_YMAX = 1024.0;

__global__ void k_pois_pack(PoisArgs a, PoisNode* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  double v = a.y[i];
  PoisNode nd;
  nd.y = v;
  nd.x = v <= PF_YMAX ? a.x[i] : NAN;
  nd.A = v > 0.0 ? (-0.5 * log(NIMB_2PI * v) + (-stirlerr(v))) + fma(-v, log(v), v) : 0.0;
  nd.g = pois_lam_index(a, i);
  nd.pad = 0;
  out[i] = nd;
}

// table-driven exp (fastmath.cuh), |eta| < 700: 12 FP64 instructions, one 8-byte look-up, four integer instructions
// (n & 63 -> table address, n & ~63 -> exponent increment: (n >> 6) << 20 = (n & ~63) * 2^14).
// The table is held once per lane, expT[j][lane] (16 KiB per CTA; the argument points at this lane's column): the 32 lanes'
// look-ups land in 32 different bank pairs whatever their indices, exactly two wavefronts -- with one shared 64-entry copy the
// data-dependent indices cost 4.8 wavefronts per look-up and made the load/store pipe the busiest unit (ncu r2k: 72 %).
static const int PF_EXPT_DOUBLES = 64 * 32;
__device__ __forceinline__ void pois_load_expT(double* dst, int tid, int nthreads) {
  for (int i = tid; i < PF_EXPT_DOUBLES; i += nthreads) dst[i] = g_mathtab.expT[i >> 5];
}
__device__ __forceinline__ double pois_exp_core(double eta, const double* __restrict__ expT, const double (&K)[7]) {
  double t = fma(eta, K[0], K[1]);
  const int n = __double2loint(t);
  t = t - K[1];
  double r = fma(t, K[2], eta);
  r = fma(t, K[3], r);
  double pl = fma(r, K[4], K[5]);
  pl = fma(pl, r, K[6]);
  pl = fma(pl, r, 0.5);
  pl = fma(pl, r, 1.0);
  const double Tj = expT[(n & 63) * 32];
  const double v = fma(Tj * r, pl, Tj);
  return __hiloint2double(__double2hiint(v) + (n & ~63) * 16384, __double2loint(v));
}
// the same with the range check folded into a running maximum of |eta|'s high word (integer pipe)
__device__ __forceinline__ double pois_exp(double eta, const double* __restrict__ expT, const double (&K)[7], unsigned& worst) {
  worst = max(worst, (unsigned)(__double2hiint(eta) & 0x7fffffff));
  return pois_exp_core(eta, expT, K);
}
// a run is regular when every |eta| < 300 (not NaN: x is NaN for y > 1024) and every lam lies in [1e-150, 1e150]: then
// mu = lam exp(eta) is a normal number and the fast form below is the density to ~4e-13 relative
static const unsigned PF_ETA_LIMIT = 0x4072c000u;   // high word of 300.0
__device__ __forceinline__ bool pois_lam_regular(double lam) { return lam >= 1e-150 && lam <= 1e150; }

__device__ __forceinline__ double pois_fast(const PoisNode& nd, double lam, double L, double b0, double b1,
                                            const double* __restrict__ expT, const double (&K)[7], unsigned& worst) {
  const double eta = fma(b1, nd.x, b0);
  const double mu = lam * pois_exp(eta, expT, K, worst);
  return fma(nd.y, L + eta, nd.A) - mu;
}

// a run that holds an irregular element, redone element by element: the fast form where it applies, the generic density
// elsewhere (rare: kept out of line, out of the hot loop)
__device__ __noinline__ double pois_exact_run(const double* __restrict__ ya, const double* __restrict__ xa, const PoisNode* sn,
                                              int k_lo, int k_hi, int t0, const double* lamp, size_t C, double b0, double b1,
                                              const double* expT) {
  double K[7];
#pragma unroll
  for (int j = 0; j < 7; j++) K[j] = c_tab_misc[j];
  double s = 0.0;
  for (int k = k_lo; k < k_hi; k++) {
    const PoisNode nd = sn[k];
    const double lam = lamp[(size_t)nd.g * C];
    unsigned w = 0;
    const double v = pois_fast(nd, lam, log(lam), b0, b1, expT, K, w);
    if (w < PF_ETA_LIMIT && pois_lam_regular(lam)) s += v;
    else s += dpois(ya[t0 + k], lam * exp(fma(b1, xa[t0 + k], b0)));
  }
  return s;
}

static const int PF_THREADS = 256;  // 8 warps x 32 chains
static const int PF_TILE = 512;     // nodes staged per step (16 KiB; with the 16 KiB exp table the static shared memory stays under 48 KiB)

// block sums: grid (chain tiles of 32, node chunks); lane = chain (lam gathers coalesced over the warp), the 8 warps of a
// CTA share the staged node tile (warp w takes the w-th run of 64 consecutive nodes, two nodes per trip so that two
// independent dependency chains are in flight).  partials[chunk][C].
// Round 2: 16 FP64 instructions per evaluation against ~14 others (round 1: 16.5 against 31.5, issue-bound) -- the per-element
// range tests and selects became one running maximum per run, the group test is per pair of nodes.
__global__ void __launch_bounds__(PF_THREADS, 3)
k_pois_sum(DevTab T, PoisArgs a, const PoisNode* __restrict__ pk, int lo, int hi, int nodes_per_chunk, double* __restrict__ partials) {
  __shared__ PoisNode sn[PF_TILE];
  __shared__ double red[PF_THREADS / 32][32];
  __shared__ double expTs[PF_EXPT_DOUBLES];
  pois_load_expT(expTs, threadIdx.x, PF_THREADS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = PF_THREADS / 32;
  const double* __restrict__ expT = expTs + lane;
  const int c = blockIdx.x * 32 + lane;
  const bool live = c < T.C;
  const int cc = live ? c : T.C - 1;
  const size_t C = (size_t)T.C;
  const double b0 = T.state[(size_t)a.b0_slot * C + cc], b1 = T.state[(size_t)a.b1_slot * C + cc];
  const double* __restrict__ lamp = T.state + (size_t)a.lam_slot * C + cc;
  const int c_lo = lo + blockIdx.y * nodes_per_chunk, c_hi = min(hi, c_lo + nodes_per_chunk);
  double K[7];
#pragma unroll
  for (int j = 0; j < 7; j++) K[j] = c_tab_misc[j];
  double acc0 = 0.0, acc1 = 0.0;
  for (int t0 = c_lo; t0 < c_hi; t0 += PF_TILE) {
    const int cnt = min(PF_TILE, c_hi - t0);
    __syncthreads();
    {  // the packed nodes of a tile are contiguous: 16-byte coalesced copies
      const int4* __restrict__ src = reinterpret_cast<const int4*>(pk + t0);
      int4* dst = reinterpret_cast<int4*>(sn);
      for (int q = threadIdx.x; q < cnt * 2; q += PF_THREADS) dst[q] = src[q];
    }
    __syncthreads();
    const int per = PF_TILE / nw, k_lo = warp * per, k_hi = min(cnt, (warp + 1) * per);
    int gprev = -1;
    double lam = 0.0, L = 0.0, r0 = 0.0, r1 = 0.0;
    unsigned worst = 0;
    auto regroup = [&](int g) {  // new group: this chain's lam and its logarithm
      gprev = g;
      lam = lamp[(size_t)g * C];
      if (!pois_lam_regular(lam)) worst = 0x7ff00000u;
      L = log(lam);
    };
    int k = k_lo;
    for (; k + 1 < k_hi; k += 2) {
      const PoisNode n0 = sn[k], n1 = sn[k + 1];
      double lam0 = lam, L0 = L;
      if ((n0.g ^ gprev) | (n1.g ^ gprev)) {   // node-level, hence warp-uniform
        if (n0.g != gprev) regroup(n0.g);
        lam0 = lam; L0 = L;
        if (n1.g != gprev) regroup(n1.g);
      }
      r0 += pois_fast(n0, lam0, L0, b0, b1, expT, K, worst);
      r1 += pois_fast(n1, lam, L, b0, b1, expT, K, worst);
    }
    if (k < k_hi) {
      const PoisNode n0 = sn[k];
      if (n0.g != gprev) regroup(n0.g);
      r0 += pois_fast(n0, lam, L, b0, b1, expT, K, worst);
    }
    if (worst >= PF_ETA_LIMIT) acc1 += pois_exact_run(a.y, a.x, sn, k_lo, k_hi, t0, lamp, C, b0, b1, expT);
    else { acc0 += r0; acc1 += r1; }
  }
  red[warp][lane] = acc0 + acc1;
  __syncthreads();
  if (warp == 0 && live) {
    double s = 0.0;
    for (int w = 0; w < nw; w++) s += red[w][lane];
    partials[(size_t)blockIdx.y * T.C + c] = s;
  }
}

// data-only statistics of each target's nodes, computed once per sampler: stats[t] = sum y (integers: exact in any order),
// stats[nT + t] = sum of the node constants A, stats[2 nT + t] = sum x y -- these two serve the block-sum by-product below --
// stats[3 nT + t] = xm = max |x| (this rank's nodes), and from stats[4 nT] on the scaled moments
//   mom[t][k] = sum_i (x_i / xm)^k / k!,  k = 0 .. PF_MOM_K,
// the coefficients of the group's moment series (k_pois_S).
static const int PF_MOM_K = 24;
static const int PF_MOM_N = PF_MOM_K + 1;
static const double PF_MOM_UMAX = 2.0;   // the series serves |b1| max|x| <= 2
__host__ __device__ inline size_t pois_stats_doubles(size_t nT) { return nT * (4 + PF_MOM_N) + 1; }   // + max over targets of xm
__global__ void k_pois_group_ysum(DevTab T, int sid, PoisArgs a, const PoisNode* __restrict__ pk, double* __restrict__ stats) {
  const nimb200_sampler sp = T.samplers[sid];
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= sp.n_targets) return;
  const nimb200_dep dp = T.deps[sp.dep_off + 1];
  Sel s = sel_of(T, dp, t);
  double c0 = 0.0, sa = 0.0, sxy = 0.0, xm = 0.0;
  for (int k = s.lo; k < s.hi; k++) {
    int node = s.idx ? s.idx[k] : k;
    c0 += a.y[node];
    sa += pk[node].A;
    sxy += a.x[node] * a.y[node];
    const double ax = fabs(a.x[node]);
    xm = ax <= xm ? xm : ax;   // a NaN sticks (and then disables the group-sum path for this target)
  }
  stats[t] = c0;
  stats[(size_t)sp.n_targets + t] = sa;
  stats[2 * (size_t)sp.n_targets + t] = sxy;
  stats[3 * (size_t)sp.n_targets + t] = xm;
  // max over the targets (non-negative doubles order like their bit patterns; a NaN counts as +Inf): the chain list's bound
  atomicMax(reinterpret_cast<unsigned long long*>(stats + pois_stats_doubles((size_t)sp.n_targets) - 1),
            (unsigned long long)__double_as_longlong(xm == xm ? xm : INFINITY));
  double mom[PF_MOM_N];
#pragma unroll
  for (int k = 0; k < PF_MOM_N; k++) mom[k] = 0.0;
  const bool usable = xm > 0.0 && xm < 1e300;   // all-zero covariates: mom[0] = n is the whole series
  for (int k = s.lo; k < s.hi; k++) {
    const int node = s.idx ? s.idx[k] : k;
    const double r = usable ? a.x[node] / xm : 0.0;
    double term = 1.0;   // r^k / k!
#pragma unroll
    for (int j = 0; j < PF_MOM_N; j++) {
      mom[j] += term;
      term = term * r / (double)(j + 1);
    }
  }
  double* out = stats + 4 * (size_t)sp.n_targets + (size_t)t * PF_MOM_N;
#pragma unroll
  for (int k = 0; k < PF_MOM_N; k++) out[k] = mom[k];
}

// conjugate gamma-Poisson sweep over the lam[g] targets (MCMC_conjugacy.R:80-93: contribution_shape = value,
// contribution_rate = coeff, coeff = the dependent's lambda with the target set to 1 = exp(b0 + b1 x)).
// thread = (target g, chain c), chains fastest: the 32 lanes of a warp read the same node (one broadcast load) and differ
// in (b0, b1).  Two nodes per trip (two independent dependency chains), range check on the exponent bits.
// PHASE 0: whole update (contrib, if given, receives each (target, chain)'s share of the dependents' block sum under the new
// value).  Data-sharded runs: PHASE 1 leaves this rank's share of the rate contribution in part[t][c]
// (sel_of clamps the target's nodes to the rank's slice), the caller all-reduces it, PHASE 2 draws from the reduced sums.
// Group sums (round 2): exp(b0 + b1 x) = exp(b0) exp(b1 x), so  S[t][c] = sum_i exp(b1 x_i)  over a target's nodes depends on b1
// alone.  PHASE 3 computes just that (part = S; the one pass over the nodes that an iteration still needs, when b1 moves);
// PHASE 4 is the whole update from a valid S (c1 = exp(b0) S: no node is read).
template <int PHASE>
__global__ void __launch_bounds__(PF_THREADS, 4)
k_pois_conj_sweep(DevTab T, int sid, uint64_t seed, uint64_t iter, PoisArgs a, const double* __restrict__ ysum, double* __restrict__ part,
                  double* __restrict__ contrib, const double* __restrict__ chainpar = nullptr) {
  __shared__ double expT[64];
  if (threadIdx.x < 64) expT[threadIdx.x] = g_mathtab.expT[threadIdx.x];
  __syncthreads();
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c;
  size_t tq;
  split_index(gid, T.C, tq, c);
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq;
  if (PHASE == 0 || PHASE == 3 || PHASE == 4) {  // targets sharded by ownership (group-aligned shards): the owner alone updates a target
    const nimb200_dep tp0 = T.deps[sp.dep_off];
    const int tn = (tp0.sel == NIMB_SEL_SAME ? tp0.a + t : tp0.a);
    if (tn < T.shard_lo[tp0.decl] || tn >= T.shard_hi[tp0.decl]) return;
  }
  const double b0s = T.state[(size_t)a.b0_slot * T.C + c], b1 = T.state[(size_t)a.b1_slot * T.C + c];
  const double b0 = PHASE == 3 ? 0.0 : b0s;
  const nimb200_dep dp = T.deps[sp.dep_off + 1];
  Sel s = sel_of(T, dp, t);
  double K[7];
#pragma unroll
  for (int j = 0; j < 7; j++) K[j] = c_tab_misc[j];
  auto expf = [&](double eta, unsigned& worst) {  // exp_tab with the range check folded into a running maximum of |eta|'s high word
    worst = max(worst, (unsigned)(__double2hiint(eta) & 0x7fffffff));
    double tt = fma(eta, K[0], K[1]);
    const int n = __double2loint(tt);
    tt = tt - K[1];
    double r = fma(tt, K[2], eta);
    r = fma(tt, K[3], r);
    double pl = fma(r, K[4], K[5]);
    pl = fma(pl, r, K[6]);
    pl = fma(pl, r, 0.5);
    pl = fma(pl, r, 1.0);
    const double Tj = expT[n & 63];
    const double v = fma(Tj * r, pl, Tj);
    return __hiloint2double(__double2hiint(v) + ((n >> 6) << 20), __double2loint(v));
  };
  double c1 = 0.0, c1b = 0.0;
  unsigned worst = 0;
  int k = s.lo;
  if (PHASE == 2) { k = s.hi; c1 = part[(size_t)t * T.C + c]; }
  if (PHASE == 4) {  // regular chains: exp(b0) S; anything else (|b0| >= 300, S marked NaN by PHASE 3) walks the nodes below
    const double Sg = part[(size_t)t * T.C + c];
    if (fabs(b0) < 300.0 && isfinite(Sg)) { k = s.hi; c1 = (chainpar ? chainpar[c] : exp(b0)) * Sg; }
    else { k = s.hi; worst = 0x7ff00000u; }
  }
  for (; k + 1 < s.hi; k += 2) {
    const int n0 = s.idx ? s.idx[k] : k, n1 = s.idx ? s.idx[k + 1] : k + 1;
    const double x0 = a.x[n0], x1 = a.x[n1];
    c1 += expf(fma(b1, x0, b0), worst);
    c1b += expf(fma(b1, x1, b0), worst);
  }
  if (k < s.hi) c1 += expf(fma(b1, a.x[s.idx ? s.idx[k] : k], b0), worst);
  c1 += c1b;
  if (PHASE != 2 && worst >= 0x4085e000u) {  // |eta| >= 700 or NaN somewhere: redo with the library exp
    c1 = 0.0;
    for (int kk = s.lo; kk < s.hi; kk++) {
      int node = s.idx ? s.idx[kk] : kk;
      c1 += 1.0 * exp(b0 + b1 * a.x[node]);
    }
  }
  // PHASE 3: a group whose |b1 x| reaches 300 is marked NaN -- its users then take their exact per-node paths, because
  // exp(b0) exp(b1 x) and exp(b0 + b1 x) part ways once either factor over- or underflows
  if (PHASE == 3) { part[(size_t)t * T.C + c] = worst >= 0x4072c000u ? NAN : c1; return; }
  if (PHASE == 1) { part[(size_t)t * T.C + c] = c1; return; }
  const double c0 = ysum[t];
  const nimb200_dep tp = T.deps[sp.dep_off];
  int tnode = (tp.sel == NIMB_SEL_SAME ? tp.a + t : tp.a);
  // chain-level quantities (exp(b0), the prior's shape and rate when their programs read nothing indexed by the node) come
  // from k_pois_chain_params: the same operations, once per chain instead of once per (target, chain)
  double prior_shape = chainpar ? chainpar[(size_t)T.C + c] : get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
  double prior_rate = chainpar ? chainpar[2 * (size_t)T.C + c] : get_param(T, tp.decl, tnode, c, NIMB_PARAM_RATE, -1, 0.0);
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double shape = prior_shape + c0, scale = 1 / (prior_rate + c1);
  const double nl = rgamma1(key, shape) * scale;
  T.state[(size_t)(sp.target_slot + t) * T.C + c] = nl;
  if ((PHASE == 0 || PHASE == 4) && contrib) {
    // by-product: the group's share of the dependents' block log-density under the value just drawn,
    //   sum_i [A_i + y_i (log lam + b0 + b1 x_i) - lam exp(b0 + b1 x_i)] = SA + Sy (log lam + b0) + b1 Sxy - lam c1,
    // so that the next MH step on this declaration finds its current-state sum without another pass over the nodes
    const double sa = ysum[(size_t)sp.n_targets + t], sxy = ysum[2 * (size_t)sp.n_targets + t];
    double v;
    if (nl == 0.0) v = c0 == 0.0 ? 0.0 : -INFINITY;        // dpois(y, 0)
    else {   // table-driven logarithm for a positive normal draw (tables read through L1: a CTA serves only 256 pairs)
      const double lg = (nl >= 1e-300 && nl < 1e300) ? log_tab(nl, &g_mathtab) : log(nl);
      v = ((sa + c0 * (lg + b0)) + b1 * sxy) - nl * c1;
    }
    contrib[(size_t)t * T.C + c] = v;
  }
}

// S[t][c] = sum over target t's nodes of exp(b1_c x_i): what an iteration of the group-sum path still computes when b1 moves.
// Moment series (round 2): with u = b1 xm,  sum_i exp(b1 x_i) = sum_k u^k mom[t][k]  -- the group's covariates enter through
// PF_MOM_N data-only coefficients, so a (target, chain) costs 24 fused multiply-adds (Horner) instead of one exp per node
// (~1200 FP64 instructions for the benchmark's 100 nodes per group).  Truncation: |u| <= 2 leaves 2^25/25! = 2e-18 of n;
// rounding: the terms are bounded by n e^|u| against a sum of at least n e^-|u|, i.e. <= e^4 x a few ulp (k_pois_S_series).
// Chains with |u| > 2 for some target are listed (k_pois_chain_list) and take the node pass (k_pois_S): grid (tiles of 32
// listed chains, blocks of PS3_TPB targets), lane = chain, warp w takes targets w, w + 8, ..; the warp stages the covariates in
// shared memory (coalesced loads, broadcast reads), four nodes per trip = four independent 12-instruction exp chains per lane
// and five other instructions per evaluation.
// A (target, chain) whose |b1| max|x| reaches 300 (or is NaN) is marked NaN -- its users then take their exact per-node paths,
// because exp(b0) exp(b1 x) and exp(b0 + b1 x) part ways once either factor over- or underflows.  Targets sharded by ownership
// (group-aligned shards): the owner alone computes a target's S.
static const int PS3_TPB = 64;
static const int PS3_XCH = 128;   // nodes of a target staged per step and warp
static const int PSS_CPT = 4;     // chains per thread of the series kernel

// the series part: lane = chain, warp w takes targets w, w + 8, ..; writes S where |b1| max|x|_t <= 2 and nothing elsewhere
// (those (target, chain) pairs belong to chains of the list below and are written by the node pass).  The block's 64 x 25
// coefficients are staged in shared memory (coalesced), then read as broadcast 16-byte loads: one load per two
// multiply-adds instead of one global load per multiply-add (0.58 ms per call at 1e4 targets x 8192 chains before, ncu r2u).
__global__ void __launch_bounds__(PF_THREADS)
k_pois_S_series(DevTab T, int sid, PoisArgs a, const double* __restrict__ stats, double* __restrict__ S) {
  __shared__ __align__(16) double smom[PS3_TPB][PF_MOM_N + 1];   // 26 doubles per target: rows stay 16-byte aligned
  const nimb200_sampler sp = T.samplers[sid];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t C = (size_t)T.C;
  const int t_lo = blockIdx.y * PS3_TPB, t_hi = min(sp.n_targets, t_lo + PS3_TPB);
  const double* __restrict__ xmax = stats + 3 * (size_t)sp.n_targets;
  const double* __restrict__ mom = stats + 4 * (size_t)sp.n_targets + (size_t)t_lo * PF_MOM_N;
  for (int q = threadIdx.x; q < (t_hi - t_lo) * PF_MOM_N; q += PF_THREADS) smom[q / PF_MOM_N][q % PF_MOM_N] = mom[q];
  __syncthreads();
  // PSS_CPT chains per thread (lane, lane + 32, ..): one broadcast read of two coefficients feeds 2 x PSS_CPT multiply-adds --
  // with one chain per thread the load/store pipe was the busiest unit (89 %, ncu r2u)
  int cs[PSS_CPT];
  double b1[PSS_CPT];
#pragma unroll
  for (int q = 0; q < PSS_CPT; q++) {
    cs[q] = blockIdx.x * (32 * PSS_CPT) + q * 32 + lane;
    b1[q] = T.state[(size_t)a.b1_slot * C + min(cs[q], T.C - 1)];
  }
  const nimb200_dep tp0 = T.deps[sp.dep_off];
  for (int t = t_lo + warp; t < t_hi; t += PF_THREADS / 32) {
    const int tn = (tp0.sel == NIMB_SEL_SAME ? tp0.a + t : tp0.a);
    if (tn < T.shard_lo[tp0.decl] || tn >= T.shard_hi[tp0.decl]) continue;
    const double xm = xmax[t];
    const double2* __restrict__ mk = reinterpret_cast<const double2*>(smom[t - t_lo]);   // the same address for all lanes: broadcast
    double u[PSS_CPT], h[PSS_CPT];
#pragma unroll
    for (int q = 0; q < PSS_CPT; q++) { u[q] = b1[q] * xm; h[q] = smom[t - t_lo][PF_MOM_K]; }
#pragma unroll
    for (int k = PF_MOM_K / 2 - 1; k >= 0; k--) {
      const double2 m2 = mk[k];
#pragma unroll
      for (int q = 0; q < PSS_CPT; q++) {
        h[q] = fma(h[q], u[q], m2.y);
        h[q] = fma(h[q], u[q], m2.x);
      }
    }
#pragma unroll
    for (int q = 0; q < PSS_CPT; q++)
      if (cs[q] < T.C && fabs(u[q]) <= PF_MOM_UMAX) S[(size_t)t * C + cs[q]] = h[q];
  }
}

// the chains that need the node pass for at least one target, in chain order: |b1| max_t max|x|_t > 2 (or NaN).  One warp,
// ballot compaction: list[0 .. count), count in list[C].  While the RW scale on b1 is unadapted that is about half of the
// chains, a few per cent of them for hundreds of iterations afterwards, none in the end -- and the node pass costs in
// proportion, where a pass by tiles of 32 consecutive chains ran in full as long as one chain per tile was left.
// One block of 1024 threads, ballots + a scan of the warp counts per 1024 chains (a single warp took 120 us for 8192 chains).
__global__ void __launch_bounds__(1024) k_pois_chain_list(DevTab T, PoisArgs a, const double* __restrict__ xmax_all, int* __restrict__ list) {
  __shared__ int wcount[32], woff[32], s_total;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double xm = *xmax_all;
  int count = 0;   // chains listed so far (the same in every thread)
  for (int base = 0; base < T.C; base += 1024) {
    const int c = base + threadIdx.x;
    const bool need = c < T.C && !(fabs(T.state[(size_t)a.b1_slot * T.C + c]) * xm <= PF_MOM_UMAX);
    const unsigned mask = __ballot_sync(0xffffffffu, need);
    if (lane == 0) wcount[warp] = __popc(mask);
    __syncthreads();
    if (warp == 0) {   // exclusive scan of the 32 warp counts
      const int v = wcount[lane];
      int incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int up = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += up; }
      woff[lane] = incl - v;
      if (lane == 31) s_total = incl;
    }
    __syncthreads();
    if (need) list[count + woff[warp] + __popc(mask & ((1u << lane) - 1u))] = c;
    count += s_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) list[T.C] = count;
}

// the node pass.  list == nullptr: every chain, tile blockIdx.x of 32 consecutive chains (no series: NIMB200_NO_SERIES);
// else: the chains of the list, tile blockIdx.x of it, and only the (target, chain) pairs the series did not serve are written.
__global__ void __launch_bounds__(PF_THREADS, 4)
k_pois_S(DevTab T, int sid, PoisArgs a, const double* __restrict__ stats, double* __restrict__ S, const int* __restrict__ list) {
  __shared__ double expTs[PF_EXPT_DOUBLES];
  __shared__ __align__(16) double xs[PF_THREADS / 32][PS3_XCH];
  const int count = list ? list[T.C] : T.C;
  if (blockIdx.x * 32 >= count) return;
  pois_load_expT(expTs, threadIdx.x, PF_THREADS);
  __syncthreads();
  const nimb200_sampler sp = T.samplers[sid];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double* __restrict__ expT = expTs + lane;
  const int ci = blockIdx.x * 32 + lane;
  const bool live = ci < count;
  const int c = list ? list[live ? ci : count - 1] : (live ? ci : count - 1);
  const size_t C = (size_t)T.C;
  const double b1 = T.state[(size_t)a.b1_slot * C + c];
  const nimb200_dep tp0 = T.deps[sp.dep_off], dp = T.deps[sp.dep_off + 1];
  const int t_lo = blockIdx.y * PS3_TPB, t_hi = min(sp.n_targets, t_lo + PS3_TPB);
  const double* __restrict__ xmax = stats + 3 * (size_t)sp.n_targets;
  double K[7];
#pragma unroll
  for (int j = 0; j < 7; j++) K[j] = c_tab_misc[j];
  const double* __restrict__ xa = a.x;
  double* __restrict__ xw = xs[warp];
  for (int t = t_lo + warp; t < t_hi; t += PF_THREADS / 32) {
    const int tn = (tp0.sel == NIMB_SEL_SAME ? tp0.a + t : tp0.a);
    if (tn < T.shard_lo[tp0.decl] || tn >= T.shard_hi[tp0.decl]) continue;
    const Sel s = sel_of(T, dp, t);
    const double u = b1 * xmax[t];
    const bool regular = fabs(u) < 300.0;
    const bool series = list != nullptr && fabs(u) <= PF_MOM_UMAX;   // served by k_pois_S_series
    const double bb = regular ? b1 : 0.0;   // keeps the table index in range; the result is discarded
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    for (int k0 = s.lo; k0 < s.hi; k0 += PS3_XCH) {
      // the warp stages the next (up to) 128 covariates of the target: coalesced loads, then broadcast 16-byte reads
      const int cnt = min(PS3_XCH, s.hi - k0);
      __syncwarp();
#pragma unroll
      for (int q = 0; q < PS3_XCH / 32; q++) {
        const int k = k0 + q * 32 + lane;
        if (k < s.hi) xw[q * 32 + lane] = __ldg(xa + (s.idx ? s.idx[k] : k));
      }
      __syncwarp();
      int k = 0;
      for (; k + 3 < cnt; k += 4) {
        const double2 xa01 = *reinterpret_cast<const double2*>(xw + k), xa23 = *reinterpret_cast<const double2*>(xw + k + 2);
        s0 += pois_exp_core(bb * xa01.x, expT, K);
        s1 += pois_exp_core(bb * xa01.y, expT, K);
        s2 += pois_exp_core(bb * xa23.x, expT, K);
        s3 += pois_exp_core(bb * xa23.y, expT, K);
      }
      for (; k < cnt; k++) s0 += pois_exp_core(bb * xw[k], expT, K);
    }
    if (live && !series) S[(size_t)t * C + c] = regular ? (s0 + s1) + (s2 + s3) : NAN;
  }
}

// per chain: exp(b0), prior shape, prior rate of the sweep's targets (chain-level parameter programs only) -> out[3][C]
__global__ void k_pois_chain_params(DevTab T, int sid, PoisArgs a, double* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= T.C) return;
  const nimb200_sampler sp = T.samplers[sid];
  const nimb200_dep tp = T.deps[sp.dep_off];
  const int tnode = tp.sel == NIMB_SEL_SAME ? tp.a : tp.a;
  out[c] = exp(T.state[(size_t)a.b0_slot * T.C + c]);
  out[(size_t)T.C + c] = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
  out[2 * (size_t)T.C + c] = get_param(T, tp.decl, tnode, c, NIMB_PARAM_RATE, -1, 0.0);
}

// Whole-declaration block sum from the data-only group sums (ysum = [sum y | sum A | sum x y] per target) and S:
//   sum_i [A_i + y_i (log lam_g + b0 + b1 x_i) - lam_g exp(b0) exp(b1 x_i)]
//     = sum_g [ SA_g + Sy_g (log lam_g + b0) + b1 Sxy_g - lam_g exp(b0) S_g ].
// grid (chain tiles of 32, target chunks), lane = chain, the 8 warps of a block stride over the chunk's targets;
// partials[chunk][C].  lam = 0 reproduces dpois(y, 0) (0 if the group's counts are all zero, -Inf otherwise); lam < 0 or NaN
// gives NaN as the generic path does.
static const int PS_TPC = 256;   // targets per chunk
__global__ void __launch_bounds__(256)
k_pois_sum_from_S(DevTab T, int sid, PoisArgs a, const double* __restrict__ ysum, const double* __restrict__ S, int nT, int t_lo,
                  int t_hi, double* __restrict__ partials) {
  __shared__ double red[8][32];
  __shared__ __align__(16) MathTab tab;
  load_mathtab(&tab, threadIdx.x, 256);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int cc = c < T.C ? c : T.C - 1;
  const size_t C = (size_t)T.C;
  const double b0 = T.state[(size_t)a.b0_slot * C + cc], b1 = T.state[(size_t)a.b1_slot * C + cc];
  const double eb0 = exp(b0);
  const double* __restrict__ lamp = T.state + (size_t)a.lam_slot * C + cc;
  const int lo = t_lo + blockIdx.y * PS_TPC, hi = min(t_hi, lo + PS_TPC);
  double acc = 0.0;
  // four targets per trip with their loads issued together (the kernel moves 16 bytes per term: memory-level parallelism is
  // what it needs); the terms are added in target order as before
  int t = lo + warp;
  for (; t + 24 < hi; t += 32) {
    double lam4[4], S4[4], sy4[4], sa4[4], sxy4[4], v4[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int tj = t + 8 * j;
      lam4[j] = lamp[(size_t)tj * C]; S4[j] = S[(size_t)tj * C + cc];
      sy4[j] = ysum[tj]; sa4[j] = ysum[(size_t)nT + tj]; sxy4[j] = ysum[2 * (size_t)nT + tj];
    }
    bool irregular = !(fabs(b0) < 300.0);
#pragma unroll
    for (int j = 0; j < 4; j++) irregular |= !isfinite(S4[j]) || !(lam4[j] >= 1e-300 && lam4[j] < 1e300);
    if (irregular) break;   // this lane finishes its targets one by one below (exact per-node path where needed)
    // table-driven logarithm (fastmath.cuh, <= 1 ulp): the library's cost 100 of the 120 instructions per term and made
    // this 16-bytes-per-term kernel compute-bound (ncu r2u: FP64 pipe 49 %, 0.40 ms for 1.3 GB)
#pragma unroll
    for (int j = 0; j < 4; j++) v4[j] = ((sa4[j] + sy4[j] * (log_tab(lam4[j], &tab) + b0)) + b1 * sxy4[j]) - lam4[j] * (eb0 * S4[j]);
#pragma unroll
    for (int j = 0; j < 4; j++) acc += v4[j];
  }
  for (; t < hi; t += 8) {
    const double sy = ysum[t], sa = ysum[(size_t)nT + t], sxy = ysum[2 * (size_t)nT + t];
    const double lam = lamp[(size_t)t * C], Sg = S[(size_t)t * C + cc];
    double v;
    if (!(fabs(b0) < 300.0) || !isfinite(Sg) || !(lam > 0.0 && lam < 1e300)) {
      // irregular (target, chain): node by node through the generic density, as the node kernel's exact path does
      const nimb200_sampler sp = T.samplers[sid];
      Sel sl = sel_of(T, T.deps[sp.dep_off + 1], t);
      v = 0.0;
      for (int k = sl.lo; k < sl.hi; k++) {
        const int node = sl.idx ? sl.idx[k] : k;
        v += dpois(a.y[node], lam * exp(fma(b1, a.x[node], b0)));
      }
    } else v = ((sa + sy * (log(lam) + b0)) + b1 * sxy) - lam * (eb0 * Sg);
    acc += v;
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < T.C) {
    double s = 0.0;
    for (int w = 0; w < 8; w++) s += red[w][lane];
    partials[(size_t)blockIdx.y * C + c] = s;
  }
}

// after the MH decision on a proposal that moved b1: the accepted chains take the proposed group sums
__global__ void k_pois_S_commit(double* __restrict__ S, const double* __restrict__ Snew, const int* __restrict__ accept, int nT, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;   // grid (chain blocks, target blocks): no 64-bit modulo per element
  if (c >= C || !accept[c]) return;
  const int t_lo = blockIdx.y * 64, t_hi = min(nT, t_lo + 64);
  int t = t_lo;
  for (; t + 7 < t_hi; t += 8) {   // eight loads in flight
    double v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = Snew[(size_t)(t + j) * C + c];
#pragma unroll
    for (int j = 0; j < 8; j++) S[(size_t)(t + j) * C + c] = v[j];
  }
  for (; t < t_hi; t++) S[(size_t)t * C + c] = Snew[(size_t)t * C + c];
}

}  // namespace nimb

// bern_logit.cuh -- log-density of  y ~ dbern(ilogit(eta))  on the table-driven exp / log of fastmath.cuh.
// Shared by the fused logistic kernel (fused_logit.cu) and the Bernoulli-logit-linear family (fused_bern.cuh).
#pragma once
#include "fastmath.cuh"

namespace nimb {

// dbinom(y, 1, ilogit(eta), log = 1) for y in {0, 1}.
// nmath: dbinom -> dbinom_raw(x, 1, p, q = 1-p):  x = 0: p < 0.1 ? -bd0(1, q) - p : log(q);  x = 1: q < 0.1 ? -bd0(1, p) - q : log(p)
// (plus p == 0 / q == 0 shortcuts).  In exact arithmetic -bd0(1, q) - p = log(q) + ((1 - q) - p), i.e. log(q) plus the rounding
// error of q = fl(1 - p), and -bd0(1, p) - q = log(p) because 1 - p is exact for p > 0.9.  So every branch is
//   log(t) + corr,  t = y ? p : q,  corr = y ? 0 : ((1 - q) - p) / q
// to < 1e-15 relative, with the same p = 1/(1 + exp(-eta)) rounding sequence as Utils.h:187 and the same limits
// (p == 0: y=0 -> 0, y=1 -> -Inf;  q == 0: y=0 -> -Inf, y=1 -> 0).  No data-dependent loop, no divergent branch.
//
// Moderate linear predictors (every |eta| of the warp's W x 32 values < 8, < 6 for the table form; i.e. all of a realistic
// regression): with a = |eta|
//   log p or log q = -log(1 + exp(-a)) - (y == (eta < 0) ? a : 0)
// exactly: no reciprocal and no correction term.  Two forms.  (1) exp and log from the tables of fastmath.cuh: 25.6 FP64
// instructions instead of 35.5 and three 8-byte look-ups -- the fused logistic kernel, whose FP64 multipliers are the saturated
// resource.  (2) SOFTPLUS: log(1 + exp(-a)) read off a table of degree-8 Taylor polynomials (centres k * 3/32): 12 FP64
// instructions, relative error <= 2.3e-16, but 80 bytes of shared-memory reads per value -- right for the latency-bound per-lane
// callers (Bernoulli-logit-linear family), wrong for the fused kernel: with realistic predictors the lanes hit ~40 different
// rows and the kernel turns shared-memory-bound (6.4 ms instead of 5.1; 4.8 only when all predictors are tiny).  For a < 8
// the reference's own value (p rounded, then log) carries <= 2e-16 absolute error on a result >= 3.4e-4 in magnitude, so the two
// agree to < 7e-13 relative (< 2e-13 for a < 6); beyond that the rounding of p = fl(1/(1+e)) matters (q = 1 - p loses bits,
// q == 0 from eta = 37) and the general path below reproduces it step by step.  The choice is a warp vote (VOTE), so the fused epilogue
// stays convergent; callers whose lanes have already diverged (VOTE = false) choose per lane.
//
// W values are carried through every step together (statement-level interleaving): the dependent DFMA chains of
// exp / reciprocal / log are long, and with 4 resident warps per scheduler the FP64 pipe only fills when each warp
// offers several independent chains (ncu round 1: "wait" was the top stall reason of the scalar version).

// exact but slow: |eta| so large that exp(-eta) overflows (eta < -690), or NaN.  Library exp / division / log.
static __device__ __noinline__ double bern_logit_lp_slow(double eta, bool y1) {
  double p = 1.0 / (1.0 + exp(-eta));
  if (p != p) return p;
  double q = 1.0 - p;
  double t = y1 ? p : q;
  double corr = (y1 || q == 0.0) ? 0.0 : ((1.0 - q) - p) / q;
  return log(t) + corr;
}

// SOFTPLUS: fast path on the softplus table sp (shared memory); else on the exp / log tables tab.  tab also serves the general path.
// y1v: the value of each of the W nodes (the fused logistic kernel's W evaluations share one node, see the wrapper below)
template <int W, bool VOTE = true, bool SOFTPLUS = false>
__device__ __forceinline__ void bern_logit_lp_vecy(const double (&eta)[W], const bool (&y1v)[W], double (&out)[W],
                                                   const double* __restrict__ sp, const MathTab* __restrict__ tab) {
  double x[W], t[W], r[W], pl[W], e[W], d[W], p[W], q[W], u[W];
  int n[W];
  {  // the moderate case is recognised on the exponent bits (integer pipe), before anything touches the FP64 pipe
    unsigned amax = 0;
#pragma unroll
    for (int j = 0; j < W; j++) amax = max(amax, (unsigned)(__double2hiint(eta[j]) & 0x7fffffff));
    const bool mod = amax < (SOFTPLUS ? 0x40180000u : 0x40200000u);  // every |eta| < 6 (table range) / < 8 (NaN fails the test)
    if (VOTE ? __all_sync(0xffffffffu, mod) : mod) {
      if (SOFTPLUS) {
#pragma unroll
        for (int j = 0; j < W; j++) x[j] = fabs(eta[j]);
#pragma unroll
        for (int j = 0; j < W; j++) t[j] = fma(x[j], 32.0 / 3.0, 6755399441055744.0);  // k = round(a * 32/3) in the low word
#pragma unroll
        for (int j = 0; j < W; j++) { n[j] = __double2loint(t[j]); t[j] = t[j] - 6755399441055744.0; }
#pragma unroll
        for (int j = 0; j < W; j++) r[j] = fma(-t[j], 0.09375, x[j]);  // a - k * 3/32 (the product is exact)
        const double2* row[W];
        double2 c8[W], c6[W], c4[W], c2[W], c0[W];
#pragma unroll
        for (int j = 0; j < W; j++) row[j] = reinterpret_cast<const double2*>(sp + n[j] * 10);
#pragma unroll
        for (int j = 0; j < W; j++) { c8[j] = row[j][4]; c6[j] = row[j][3]; }
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(c8[j].x, r[j], c6[j].y);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c6[j].x);
#pragma unroll
        for (int j = 0; j < W; j++) c4[j] = row[j][2];
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c4[j].y);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c4[j].x);
#pragma unroll
        for (int j = 0; j < W; j++) c2[j] = row[j][1];
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c2[j].y);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c2[j].x);
#pragma unroll
        for (int j = 0; j < W; j++) c0[j] = row[j][0];
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c0[j].y);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c0[j].x);
#pragma unroll
        for (int j = 0; j < W; j++) {
          bool neg = __double2hiint(eta[j]) < 0;
          out[j] = (y1v[j] == neg ? -x[j] : 0.0) - pl[j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < W; j++) x[j] = __hiloint2double(__double2hiint(eta[j]) | (int)0x80000000, __double2loint(eta[j]));  // -|eta|
#pragma unroll
        for (int j = 0; j < W; j++) t[j] = fma(x[j], c_tab_misc[0], c_tab_misc[1]);
#pragma unroll
        for (int j = 0; j < W; j++) { n[j] = __double2loint(t[j]); t[j] = t[j] - c_tab_misc[1]; }
#pragma unroll
        for (int j = 0; j < W; j++) r[j] = fma(t[j], c_tab_misc[2], x[j]);
#pragma unroll
        for (int j = 0; j < W; j++) r[j] = fma(t[j], c_tab_misc[3], r[j]);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(r[j], c_tab_misc[4], c_tab_misc[5]);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c_tab_misc[6]);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], 0.5);
#pragma unroll
        for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], 1.0);
#pragma unroll
        for (int j = 0; j < W; j++) {
          double T = tab->expT[n[j] & 63];
          double v = fma(T * r[j], pl[j], T);
          e[j] = __hiloint2double(__double2hiint(v) + ((n[j] >> 6) << 20), __double2loint(v));
        }
#pragma unroll
        for (int j = 0; j < W; j++) d[j] = 1.0 + e[j];  // in (1, 2]
#pragma unroll
        for (int j = 0; j < W; j++) {  // log(d), d in (1, 2]: centre c = 1 + idx/128; idx = 128 (d >= 2 - 1/256) has c = 2:
          // invc[128] = 0.5 and the logarithm of the centre is ln 2 -- no exponent arithmetic on this path
          const int idx = (__double2hiint(d[j]) - 0x3ff00000 + 0x1000) >> 13;
          double rr = fma(d[j], tab->invc[idx], -1.0);
          double Q = fma(rr, c_tab_misc[8], c_tab_misc[9]);
          Q = fma(Q, rr, c_tab_misc[10]);
          Q = fma(Q, rr, c_tab_misc[11]);
          Q = fma(Q, rr, 1.0 / 3.0);
          Q = fma(Q, rr, -0.5);
          double lp = fma(rr * rr, Q, rr);
          double lg = (idx == 128 ? c_tab_misc[7] : tab->logA[idx]) + lp;
          bool neg = __double2hiint(eta[j]) < 0;
          out[j] = (y1v[j] == neg ? x[j] : 0.0) - lg;  // x = -|eta|
        }
      }
      return;
    }
  }
  bool slow = false;
#pragma unroll
  for (int j = 0; j < W; j++) slow |= !(eta[j] >= -690.0);
  if (VOTE ? __any_sync(0xffffffffu, slow) : slow) {  // warp-uniform with VOTE, practically never taken
#pragma unroll
    for (int j = 0; j < W; j++) out[j] = bern_logit_lp_slow(eta[j], y1v[j]);
    return;
  }
  // ---- e = exp(-eta): x = k ln2/64 + r, e = 2^(k>>6) T[k&63] (1 + r P(r)).  x is clamped at -708: below that e only has to
  // vanish against 1, and 2^(k>>6) can then always be applied by one integer add to the exponent field.
#pragma unroll
  for (int j = 0; j < W; j++) x[j] = fmax(-eta[j], -708.0);
#pragma unroll
  for (int j = 0; j < W; j++) t[j] = fma(x[j], c_tab_misc[0], c_tab_misc[1]);
#pragma unroll
  for (int j = 0; j < W; j++) { n[j] = __double2loint(t[j]); t[j] = t[j] - c_tab_misc[1]; }
#pragma unroll
  for (int j = 0; j < W; j++) r[j] = fma(t[j], c_tab_misc[2], x[j]);
#pragma unroll
  for (int j = 0; j < W; j++) r[j] = fma(t[j], c_tab_misc[3], r[j]);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(r[j], c_tab_misc[4], c_tab_misc[5]);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], c_tab_misc[6]);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], 0.5);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], 1.0);
#pragma unroll
  for (int j = 0; j < W; j++) {
    double T = tab->expT[n[j] & 63];
    double v = fma(T * r[j], pl[j], T);
    e[j] = __hiloint2double(__double2hiint(v) + ((n[j] >> 6) << 20), __double2loint(v));
  }
  // ---- p = 1/(1+e): d in [1, 1e300]; Newton on a >= 20-bit seed, last step on the residual
#pragma unroll
  for (int j = 0; j < W; j++) d[j] = 1.0 + e[j];
#pragma unroll
  for (int j = 0; j < W; j++) p[j] = rcp_approx(d[j]);
#pragma unroll
  for (int it = 0; it < 3; it++) {
#pragma unroll
    for (int j = 0; j < W; j++) u[j] = fma(-d[j], p[j], 1.0);
#pragma unroll
    for (int j = 0; j < W; j++) p[j] = fma(p[j], u[j], p[j]);
  }
  // ---- t = y ? p : q.  For y = 0 the rounding error delta of q = fl(1-p) is added back inside the logarithm:
  // log(q) + delta/q = log(q + delta) to first order.
#pragma unroll
  for (int j = 0; j < W; j++) q[j] = 1.0 - p[j];
  double dl[W];
#pragma unroll
  for (int j = 0; j < W; j++) {
    dl[j] = y1v[j] ? 0.0 : (1.0 - q[j]) - p[j];  // exact
    t[j] = y1v[j] ? p[j] : q[j];
  }
  // ---- log(t), t = 0 or a normal number in [2^-1000, 1]: t = 2^e m, m in [1,2), centre c = 1 + j/128 (j = 128 -> e+1, c = 1)
#pragma unroll
  for (int j = 0; j < W; j++) {
    int hi = __double2hiint(t[j]), lo = __double2loint(t[j]);
    bool zero = (hi | lo) == 0;
    int ex = (hi >> 20) - 1023;
    int fr = hi & 0x000fffff;
    int idx = (fr + 0x1000) >> 13;
    ex += idx >> 7;
    double m = __hiloint2double(fr | 0x3ff00000, lo);
    double ic = tab->invc[idx];
    bool em1 = ex == -1;
    double lc = em1 ? tab->logB[idx] : tab->logA[idx];
    double kd = em1 ? 0.0 : (double)ex;
    double rr = fma(m, ic, -1.0);
    rr = fma(dl[j] * __hiloint2double((1023 - (hi >> 20) + 1023) << 20, 0), ic, rr);  // + delta 2^-e / c
    double Q = fma(rr, c_tab_misc[8], c_tab_misc[9]);
    Q = fma(Q, rr, c_tab_misc[10]);
    Q = fma(Q, rr, c_tab_misc[11]);
    Q = fma(Q, rr, 1.0 / 3.0);
    Q = fma(Q, rr, -0.5);
    double lp = fma(rr * rr, Q, rr);
    double lg = fma(kd, c_tab_misc[7], lc + lp);
    out[j] = zero ? -INFINITY : lg;
  }
}

template <int W, bool VOTE = true, bool SOFTPLUS = false>
__device__ __forceinline__ void bern_logit_lp_vec(const double (&eta)[W], bool y1, double (&out)[W], const double* __restrict__ sp,
                                                  const MathTab* __restrict__ tab) {
  bool yv[W];
#pragma unroll
  for (int j = 0; j < W; j++) yv[j] = y1;
  bern_logit_lp_vecy<W, VOTE, SOFTPLUS>(eta, yv, out, sp, tab);
}

// ---- the moderate path on replicated tables (round 2) ----------------------------------------------------------------
// The look-up tables (2^(j/64), 1/c_j, log c_j) held 16 times, T[j][copy]: lane l reads copy l & 15, so the 32 lanes of a load
// hit 16 different bank pairs whatever their indices -- exactly the two wavefronts a 64-bit load needs (one shared copy: 4.8
// wavefronts per look-up with realistic predictors, ncu r2o).  logc[128] = ln 2 here (the centre 2 of the moderate path), so
// the look-up needs no special case; the exponent and table-index arithmetic is the four-instruction form of fused_pois.cuh.
static const int BL_REP_ENTRIES = 64 + 129 + 129;
static const int BL_REP_BYTES = BL_REP_ENTRIES * 16 * 8;
__device__ __forceinline__ void load_bern_rep_tables(double* dst, int tid, int nthreads) {
  for (int i = tid; i < BL_REP_ENTRIES * 16; i += nthreads) {
    const int e = i >> 4;
    dst[i] = e < 64 ? g_mathtab.expT[e] : e < 64 + 129 ? g_mathtab.invc[e - 64] : (e - 193 == 128 ? c_tab_misc[7] : g_mathtab.logA[e - 193]);
  }
}
// W nodes at once; expL = the replicated tables + (lane & 15); tab = the plain tables for the general path.
// is every |eta| < 8 (NaN fails)?  On the exponent bits, integer pipe.
template <int W>
__device__ __forceinline__ bool bern_logit_moderate(const double (&eta)[W]) {
  unsigned amax = 0;
#pragma unroll
  for (int j = 0; j < W; j++) amax = max(amax, (unsigned)(__double2hiint(eta[j]) & 0x7fffffff));
  return amax < 0x40200000u;
}
template <int W, int REP>
__device__ __forceinline__ void bern_logit_lp_rep_mod(const double (&eta)[W], const bool (&y1v)[W], double (&out)[W],
                                                      const double* __restrict__ expL);
template <int W, bool VOTE, int REP = 16>
__device__ __forceinline__ void bern_logit_lp_rep(const double (&eta)[W], const bool (&y1v)[W], double (&out)[W],
                                                  const double* __restrict__ expL, const MathTab* __restrict__ tab) {
  const bool mod = bern_logit_moderate<W>(eta);
  if (!(VOTE ? __all_sync(0xffffffffu, mod) : mod)) {
    bern_logit_lp_vecy<W, VOTE, false>(eta, y1v, out, nullptr, tab);
    return;
  }
  bern_logit_lp_rep_mod<W, REP>(eta, y1v, out, expL);
}
// the moderate path itself: the caller has established that every |eta| < 8
template <int W, int REP>
__device__ __forceinline__ void bern_logit_lp_rep_mod(const double (&eta)[W], const bool (&y1v)[W], double (&out)[W],
                                                      const double* __restrict__ expL) {
  const double* __restrict__ K = c_tab_misc;
  const double* __restrict__ invL = expL + 64 * REP;
  const double* __restrict__ logL = expL + (64 + 129) * REP;
  double x[W], t[W], r[W], pl[W], d[W];
  int n[W];
#pragma unroll
  for (int j = 0; j < W; j++) x[j] = __hiloint2double(__double2hiint(eta[j]) | (int)0x80000000, __double2loint(eta[j]));  // -|eta|
#pragma unroll
  for (int j = 0; j < W; j++) t[j] = fma(x[j], K[0], K[1]);
#pragma unroll
  for (int j = 0; j < W; j++) { n[j] = __double2loint(t[j]); t[j] = t[j] - K[1]; }
#pragma unroll
  for (int j = 0; j < W; j++) r[j] = fma(t[j], K[2], x[j]);
#pragma unroll
  for (int j = 0; j < W; j++) r[j] = fma(t[j], K[3], r[j]);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(r[j], K[4], K[5]);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], K[6]);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], 0.5);
#pragma unroll
  for (int j = 0; j < W; j++) pl[j] = fma(pl[j], r[j], 1.0);
#pragma unroll
  for (int j = 0; j < W; j++) {
    const double Tj = expL[(n[j] & 63) * REP];
    const double v = fma(Tj * r[j], pl[j], Tj);
    const double e = __hiloint2double(__double2hiint(v) + (n[j] & ~63) * 16384, __double2loint(v));
    d[j] = 1.0 + e;  // in (1, 2]
  }
#pragma unroll
  for (int j = 0; j < W; j++) {  // log(d): centre c = 1 + idx/128, idx = 128 is the centre 2
    const int idx = (__double2hiint(d[j]) - 0x3ff00000 + 0x1000) >> 13;
    const double rr = fma(d[j], invL[idx * REP], -1.0);
    double Q = fma(rr, K[8], K[9]);
    Q = fma(Q, rr, K[10]);
    Q = fma(Q, rr, K[11]);
    Q = fma(Q, rr, 1.0 / 3.0);
    Q = fma(Q, rr, -0.5);
    const double lp = fma(rr * rr, Q, rr);
    const double lg = logL[idx * REP] + lp;
    const bool neg = __double2hiint(eta[j]) < 0;
    out[j] = (y1v[j] == neg ? x[j] : 0.0) - lg;  // x = -|eta|
  }
}

__device__ __forceinline__ double bern_logit_lp(double eta, bool y1) {  // scalar form (diagnostics)
  double in[1] = {eta}, o[1];
  bern_logit_lp_vec<1, true, false>(in, y1, o, nullptr, &g_mathtab);
  return o[0];
}


// per-lane form without the warp vote, for callers whose lanes have already diverged
__device__ __forceinline__ double bern_logit_lp1(double eta, bool y1, const double* __restrict__ sp) {
  if (!(eta >= -690.0)) return bern_logit_lp_slow(eta, y1);
  double in[1] = {eta}, o[1];
  bern_logit_lp_vec<1, false, true>(in, y1, o, sp, &g_mathtab);
  return o[0];
}

}  // namespace nimb

// fastmath.cuh -- FP64 exp / log / reciprocal for the fused kernels (sm_100a).
//
// Why not CUDA's exp()/log(): in the fused logistic kernel they were ~2/3 of the issued instructions without being
// FP64 work -- nvcc materialises every 64-bit polynomial coefficient with a UMOV / IMAD.MOV pair at each use (ncu,
// profiles/r1a_summary.md: 198 warp instructions per 32 evaluations, 61 of them FP64).  Here the coefficients sit in
// __constant__ memory and enter the DFMAs as constant-bank operands, the range reduction is the minimum integer work,
// and there is no slow-path call.  Accuracy target: <= 1 ulp-ish (checked against libm on the GPU by
// tests/test_gpu_parity.py::test_fastmath_accuracy), far inside the 1e-12 parity bar.
#pragma once
#include <cstdint>

namespace nimb {

// exp: r^k/k! for k = 13 .. 2 (Horner order), |r| <= ln2/2  ->  truncation < 4e-18 relative
static __device__ __constant__ double c_exp_poly[12] = {
    1.0 / 6227020800.0, 1.0 / 479001600.0, 1.0 / 39916800.0, 1.0 / 3628800.0, 1.0 / 362880.0, 1.0 / 40320.0,
    1.0 / 5040.0,       1.0 / 720.0,       1.0 / 120.0,      1.0 / 24.0,      1.0 / 6.0,      0.5};
// log: 2*atanh(s) = 2s (1 + z/3 + z^2/5 + ...), z = s^2 <= 0.0295: 1/(2j+1) for j = 10 .. 1
static __device__ __constant__ double c_log_poly[10] = {1.0 / 21.0, 1.0 / 19.0, 1.0 / 17.0, 1.0 / 15.0, 1.0 / 13.0,
                                                 1.0 / 11.0, 1.0 / 9.0,  1.0 / 7.0,  1.0 / 5.0,  1.0 / 3.0};
static __device__ __constant__ double c_misc[6] = {
    1.4426950408889634,       // log2(e)
    6755399441055744.0,       // 2^52 + 2^51: round-to-integer magic
    -6.93147180559945286e-01, // -ln2_hi
    -2.31904681384629956e-17, // -ln2_lo
    6.93147180559945286e-01,  // ln2_hi
    2.31904681384629956e-17   // ln2_lo
};

__device__ __forceinline__ double rcp_approx(double a) {  // >= 20 good bits (MUFU.RCP64H)
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  return y;
}

// 1/a for normal a in [2^-1000, 2^1000]: three Newton steps, last one on the residual (faithfully rounded)
__device__ __forceinline__ double rcp_f64(double a) {
  double y = rcp_approx(a);
  double e = fma(-a, y, 1.0);
  y = fma(y, e, y);
  e = fma(-a, y, 1.0);
  y = fma(y, e, y);
  e = fma(-a, y, 1.0);
  return fma(y, e, y);
}

__device__ __forceinline__ double exp_f64(double x) {
  x = fmin(fmax(x, -746.0), 710.0);  // beyond: 0 / +Inf after scaling anyway (NaN passes through fmin/fmax? no: handled by caller)
  double t = fma(x, c_misc[0], c_misc[1]);
  int n = __double2loint(t);
  t = t - c_misc[1];
  double r = fma(t, c_misc[2], x);
  r = fma(t, c_misc[3], r);
  double p = c_exp_poly[0];
#pragma unroll
  for (int j = 1; j < 12; j++) p = fma(p, r, c_exp_poly[j]);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  int n1 = n >> 1, n2 = n - n1;
  double s1 = __hiloint2double((n1 + 1023) << 20, 0), s2 = __hiloint2double((n2 + 1023) << 20, 0);
  return (p * s1) * s2;
}

// log(t) for t >= 0 (t = 0 -> -Inf, +Inf -> +Inf, NaN -> NaN)
__device__ __forceinline__ double log_f64(double t) {
  int hi = __double2hiint(t), lo = __double2loint(t);
  int k = 0;
  if (hi < 0x00100000) {  // zero, subnormal (or negative: not produced by the callers)
    if (t == 0.0) return -INFINITY;
    t = t * 18014398509481984.0;  // 2^54
    hi = __double2hiint(t); lo = __double2loint(t);
    k = -54;
  }
  if (hi >= 0x7ff00000) return t + t;  // Inf / NaN
  k += (hi >> 20) - 1023;
  hi = (hi & 0x000fffff) | 0x3ff00000;
  if (hi > 0x3ff6a09e) { hi -= 0x00100000; k += 1; }  // m > sqrt(2): halve
  double m = __hiloint2double(hi, lo);
  double f = m - 1.0, g = m + 1.0;
  double y = rcp_approx(g);
  double e = fma(-g, y, 1.0);
  y = fma(y, e, y);
  e = fma(-g, y, 1.0);
  y = fma(y, e, y);
  double s = f * y;
  s = fma(fma(-s, g, f), y, s);  // one residual correction of the quotient
  double z = s * s;
  double P = c_log_poly[0];
#pragma unroll
  for (int j = 1; j < 10; j++) P = fma(P, z, c_log_poly[j]);
  double s2 = s + s;
  double lm = fma(s2 * z, P, s2);  // 2s + 2s z P
  double kd = (double)k;
  return fma(kd, c_misc[4], fma(kd, c_misc[5], lm));
}

// ---- table-driven exp / log for the fused epilogue ------------------------------------------------------------
// The FP64 multipliers are the saturated resource of the fused kernel (DMMA and DFMA share them), the load/store
// units are idle: range reduction against small tables held in shared memory moves work from the former to the
// latter.  exp: 64-entry 2^(j/64) table + degree-5 polynomial (12 FP64 instructions instead of 19); log: 129-entry
// table of reciprocals / logarithms of the interval centres 1 + j/128 + degree-7 log1p (10 instead of 24, and no
// reciprocal).  Tables: gen_math_tables.py (60-digit arithmetic).  <= 1 ulp; log keeps full relative accuracy as
// t -> 1 (centre 1 has invc = 1, logc = 0; the binade [0.5, 1) has its own table with ln 2 already subtracted).
#include "math_tables.inc"
struct MathTab {
  double expT[64];
  double logA[129];
  double logB[129];
  double invc[129];
};
static __device__ const MathTab g_mathtab = {NIMB_EXPT_INIT, NIMB_LOGA_INIT, NIMB_LOGB_INIT, NIMB_INVC_INIT};
// softplus table: log(1 + exp(-a)) on [0, 6] as degree-8 Taylor polynomials about the centres k * 3/32, k = 0..64 (rows of 10
// doubles: 9 coefficients + padding, so a row is five 16-byte loads).  |a - centre| <= 3/64: remainder < 1e-17.
static const int SOFTPLUS_DOUBLES = NIMB_SOFTPLUS_ROWS * 10;
struct SoftplusTab { double c[SOFTPLUS_DOUBLES]; };
static __device__ __align__(16) const SoftplusTab g_softplus = {NIMB_SOFTPLUS_INIT};
__device__ __forceinline__ void load_softplus(double* dst, int tid, int nthreads) {
  for (int i = tid; i < SOFTPLUS_DOUBLES; i += nthreads) dst[i] = g_softplus.c[i];
}
static __device__ __constant__ double c_tab_misc[12] = {
    NIMB_EXP_L, 6755399441055744.0, -NIMB_EXP_HI, -NIMB_EXP_LO,
    1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0,                              // exp(r) = 1 + r (1 + r/2 + r^2/6 + r^3/24 + r^4/120)
    0.6931471805599453,                                              // ln 2
    1.0 / 7.0, -1.0 / 6.0, 1.0 / 5.0, -1.0 / 4.0};                   // log1p(r) = r + r^2 (-1/2 + r/3 - r^2/4 + ... + r^5/7)

// scalar forms of the table-driven functions (tab: the tables, normally a shared-memory copy)
__device__ __forceinline__ double exp_tab(double x, const MathTab* __restrict__ tab) {  // x in [-708, 709]
  double t = fma(x, c_tab_misc[0], c_tab_misc[1]);
  int n = __double2loint(t);
  t = t - c_tab_misc[1];
  double r = fma(t, c_tab_misc[2], x);
  r = fma(t, c_tab_misc[3], r);
  double pl = fma(r, c_tab_misc[4], c_tab_misc[5]);
  pl = fma(pl, r, c_tab_misc[6]);
  pl = fma(pl, r, 0.5);
  pl = fma(pl, r, 1.0);
  double T = tab->expT[n & 63];
  double v = fma(T * r, pl, T);
  return __hiloint2double(__double2hiint(v) + ((n >> 6) << 20), __double2loint(v));
}
__device__ __forceinline__ double log_tab(double t, const MathTab* __restrict__ tab) {  // t: positive normal number
  int hi = __double2hiint(t), lo = __double2loint(t);
  int ex = (hi >> 20) - 1023;
  int fr = hi & 0x000fffff;
  int idx = (fr + 0x1000) >> 13;
  ex += idx >> 7;
  double m = __hiloint2double(fr | 0x3ff00000, lo);
  bool em1 = ex == -1;
  double lc = em1 ? tab->logB[idx] : tab->logA[idx];
  double kd = em1 ? 0.0 : (double)ex;
  double rr = fma(m, tab->invc[idx], -1.0);
  double Q = fma(rr, c_tab_misc[8], c_tab_misc[9]);
  Q = fma(Q, rr, c_tab_misc[10]);
  Q = fma(Q, rr, c_tab_misc[11]);
  Q = fma(Q, rr, 1.0 / 3.0);
  Q = fma(Q, rr, -0.5);
  double lp = fma(rr * rr, Q, rr);
  return fma(kd, c_tab_misc[7], lc + lp);
}
__device__ __forceinline__ void load_mathtab(MathTab* dst, int tid, int nthreads) {
  for (int i = tid; i < (int)(sizeof(MathTab) / sizeof(double)); i += nthreads) ((double*)dst)[i] = ((const double*)&g_mathtab)[i];
}

}  // namespace nimb

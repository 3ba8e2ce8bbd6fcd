// draws_dev.cuh -- counter-based random draws (DESIGN.md "draws"): Philox4x32-10 keyed by (seed, chain),
// counter (iteration, sampler, target, call).  Replaces the reference's use of R's global RNG
// (unif_rand / norm_rand, P/inst/CppCode/Utils.cpp:61, dists.cpp:954,1155; rgamma in the generated conjugate
// samplers, P/R/MCMC_conjugacy.R:1100): one independent, reproducible stream per chain, no shared state.
//
// Every floating-point step is written with round-to-nearest intrinsics (__dadd_rn, __dmul_rn, __fma_rn,
// __ddiv_rn, __dsqrt_rn) so the compiler can neither fuse nor reorder them: a host implementation that
// performs the same IEEE operations produces bit-identical normals, uniforms and gamma variates.
#pragma once
#include <cstdint>

namespace nimb {

struct DrawKey {
  uint64_t seed;
  uint32_t chain;    // global chain index
  uint64_t iter;
  uint32_t sampler;
  uint32_t target;
};

__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}

__device__ __forceinline__ double bits2unif(uint32_t lo, uint32_t hi) {
  uint64_t m = (((uint64_t)hi << 32) | lo) >> 12;
  return __dmul_rn(__dadd_rn((double)m, 0.5), 2.220446049250313e-16);
}

// Polynomial coefficients in constant memory: as literals every 64-bit coefficient is materialised by a UMOV pair at each
// use (157 of the 1035 instructions per (target, chain) of the gamma-Poisson sweep, ncu r2n); from the constant bank two of
// them arrive per LDCU.128 or directly as a DFMA operand.  The values, the operations and their order are unchanged.
static __device__ __constant__ double c_det_log[12] = {1.0 / 23.0, 1.0 / 21.0, 1.0 / 19.0, 1.0 / 17.0, 1.0 / 15.0, 1.0 / 13.0,
                                                      1.0 / 11.0, 1.0 / 9.0,  1.0 / 7.0,  1.0 / 5.0,  1.0 / 3.0,  1.0};
static __device__ __constant__ double c_det_exp[15] = {
    1.0 / 87178291200.0, 1.0 / 6227020800.0, 1.0 / 479001600.0, 1.0 / 39916800.0, 1.0 / 3628800.0, 1.0 / 362880.0, 1.0 / 40320.0,
    1.0 / 5040.0,        1.0 / 720.0,        1.0 / 120.0,       1.0 / 24.0,       1.0 / 6.0,       0.5,            1.0,          1.0};
static __device__ __constant__ double c_det_sin[9] = {1.0 / 355687428096000.0, -1.0 / 1307674368000.0, 1.0 / 6227020800.0,
                                                     -1.0 / 39916800.0,       1.0 / 362880.0,         -1.0 / 5040.0,
                                                     1.0 / 120.0,             -1.0 / 6.0,             1.0};
static __device__ __constant__ double c_det_cos[10] = {-1.0 / 6402373705728000.0, 1.0 / 20922789888000.0, -1.0 / 87178291200.0,
                                                      1.0 / 479001600.0,         -1.0 / 3628800.0,       1.0 / 40320.0,
                                                      -1.0 / 720.0,              1.0 / 24.0,             -0.5,          1.0};

__device__ __forceinline__ void unif2(const DrawKey& k, uint32_t call, double& u1, double& u2) {
  uint32_t c[4];
  c[0] = (uint32_t)k.iter;
  c[1] = (k.sampler & 0xffffu) | ((uint32_t)((k.iter >> 32) & 0xffffu) << 16);
  c[2] = k.target;
  c[3] = call;
  philox4x32_10(c, (uint32_t)k.seed, (uint32_t)(k.seed >> 32) ^ k.chain);
  u1 = bits2unif(c[0], c[1]);
  u2 = bits2unif(c[2], c[3]);
}

__device__ __forceinline__ double det_log(double x) {
  uint64_t b = (uint64_t)__double_as_longlong(x);
  int e = (int)(b >> 52) - 1023;
  double m = __longlong_as_double((long long)((b & 0xFFFFFFFFFFFFFull) | 0x3FF0000000000000ull));
  if (m > 1.4142135623730951) { m = __dmul_rn(m, 0.5); e += 1; }
  double f = __dadd_rn(m, -1.0);
  double s = __ddiv_rn(f, __dadd_rn(m, 1.0));
  double z = __dmul_rn(s, s);
  double p = c_det_log[0];
#pragma unroll
  for (int j = 1; j < 12; j++) p = __fma_rn(p, z, c_det_log[j]);
  double t = __dmul_rn(s, p);
  return __fma_rn((double)e, 0.6931471805599453, __dmul_rn(2.0, t));
}

__device__ __forceinline__ double det_exp(double x) {
  if (x < -700.0) return 0.0;
  double kd = rint(__dmul_rn(x, 1.4426950408889634));
  double r = __fma_rn(-kd, 6.93147180369123816490e-01, x);
  r = __fma_rn(-kd, 1.90821492927058770002e-10, r);
  double p = c_det_exp[0];
#pragma unroll
  for (int j = 1; j < 15; j++) p = __fma_rn(p, r, c_det_exp[j]);
  return __dmul_rn(p, __longlong_as_double((long long)((uint64_t)((long long)kd + 1023) << 52)));
}

__device__ __forceinline__ void det_sincos2pi(double u, double& s, double& c) {
  double a = __dmul_rn(u, 8.0);
  int o = (int)a;
  double f = __dadd_rn(a, -(double)o);
  double th = __dmul_rn(f, 0.7853981633974483);
  double z = __dmul_rn(th, th);
  double ps = c_det_sin[0];
#pragma unroll
  for (int j = 1; j < 9; j++) ps = __fma_rn(ps, z, c_det_sin[j]);
  double sn = __dmul_rn(th, ps);
  double pc = c_det_cos[0];
#pragma unroll
  for (int j = 1; j < 10; j++) pc = __fma_rn(pc, z, c_det_cos[j]);
  double cs = pc;
  if (o & 1) {
    double s2 = __dmul_rn(__dadd_rn(sn, cs), 0.7071067811865476);
    double c2 = __dmul_rn(__dadd_rn(cs, -sn), 0.7071067811865476);
    sn = s2; cs = c2;
  }
  double t;
  switch (o >> 1) {
    case 0: break;
    case 1: t = sn; sn = cs; cs = -t; break;
    case 2: sn = -sn; cs = -cs; break;
    default: t = sn; sn = -cs; cs = t; break;
  }
  s = sn; c = cs;
}

__device__ __forceinline__ void normal2(const DrawKey& k, uint32_t call, double& z0, double& z1) {
  double u1, u2, s, c;
  unif2(k, call, u1, u2);
  double r = __dsqrt_rn(__dmul_rn(-2.0, det_log(u1)));
  det_sincos2pi(u2, s, c);
  z0 = __dmul_rn(r, c);
  z1 = __dmul_rn(r, s);
}

// Marsaglia-Tsang gamma(a, 1): call 0 = boost uniform (a < 1), trip m: normal = call 1+2m, uniform = call 2+2m.
__device__ __forceinline__ double rgamma1(const DrawKey& k, double a, uint32_t base = 0) {
  double boost = 1.0, u, dummy, x, v;
  if (!(a > 0.0) || !isfinite(a)) return NAN;
  if (a < 1.0) {
    unif2(k, base, u, dummy);
    boost = det_exp(__ddiv_rn(det_log(u), a));
    a = __dadd_rn(a, 1.0);
  }
  double d = __dadd_rn(a, -(1.0 / 3.0));
  double c = __ddiv_rn(1.0 / 3.0, __dsqrt_rn(d));
  v = 1.0;
  for (int m = 0; m < 64; m++) {
    normal2(k, base + 1 + 2 * m, x, dummy);
    v = __dadd_rn(1.0, __dmul_rn(c, x));
    if (v <= 0.0) continue;
    v = __dmul_rn(__dmul_rn(v, v), v);
    unif2(k, base + 2 + 2 * m, u, dummy);
    double x2 = __dmul_rn(x, x);
    if (u < __dadd_rn(1.0, -__dmul_rn(__dmul_rn(0.0331, x2), x2))) break;
    double lhs = det_log(u);
    double rhs = __dadd_rn(__dmul_rn(0.5, x2), __dmul_rn(d, __dadd_rn(__dadd_rn(1.0, -v), det_log(v))));
    if (lhs < rhs) break;
  }
  return __dmul_rn(__dmul_rn(d, v), boost);
}

// beta(a, b) = X / (X + Y), X ~ gamma(a) on calls 0.., Y ~ gamma(b) on calls 256.. of the same counter
__device__ __forceinline__ double rbeta1(const DrawKey& k, double a, double b) {
  double x = rgamma1(k, a, 0), y = rgamma1(k, b, 256);
  return __ddiv_rn(x, __dadd_rn(x, y));
}

// Poisson(mu) by inversion from zero: p_0 = exp(-mu), p_{k+1} = p_k mu / (k + 1), one uniform (call `base`).  mu <= 500 (beyond it
// exp(-mu) leaves too few bits and the walk too long): larger or invalid means return NaN.
__device__ __forceinline__ double rpois_inv(const DrawKey& k, double mu, uint32_t base = 0) {
  if (!(mu >= 0.0) || mu > 500.0) return NAN;
  if (mu == 0.0) return 0.0;
  double u, dummy;
  unif2(k, base, u, dummy);
  double p = det_exp(-mu), cdf = p;
  int x = 0;
  while (u > cdf && x < 5000) { x++; p = __ddiv_rn(__dmul_rn(p, mu), (double)x); cdf = __dadd_rn(cdf, p); }
  return (double)x;
}
// binomial(n, pr) by inversion from zero: p_0 = (1 - pr)^n, p_{k+1} = p_k (n - k) / (k + 1) pr / (1 - pr), one uniform.
// n an integer <= 100000 and (1 - pr)^n not underflowing; otherwise NaN.
__device__ __forceinline__ double rbinom_inv(const DrawKey& k, double n, double pr, uint32_t base = 0) {
  if (!(n >= 0.0) || n > 100000.0 || n != floor(n) || !(pr >= 0.0) || !(pr <= 1.0)) return NAN;
  if (pr == 0.0 || n == 0.0) return 0.0;
  if (pr == 1.0) return n;
  double u, dummy;
  unif2(k, base, u, dummy);
  const double q = __dadd_rn(1.0, -pr), r = __ddiv_rn(pr, q);
  double p = det_exp(__dmul_rn(n, det_log(q)));
  if (!(p > 0.0)) return NAN;
  double cdf = p;
  int x = 0;
  const int ni = (int)n;
  while (u > cdf && x < ni) {
    p = __dmul_rn(__ddiv_rn(__dmul_rn(p, (double)(ni - x)), (double)(x + 1)), r);
    x++;
    cdf = __dadd_rn(cdf, p);
  }
  return (double)x;
}

}  // namespace nimb

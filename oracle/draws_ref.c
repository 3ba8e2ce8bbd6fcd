/*
 * draws_ref.c -- TEST INFRASTRUCTURE (parity oracle).  Independent CPU implementation of the
 * counter-based draw specification in DESIGN.md ("draws").  The reference's own RNG (R's Mersenne
 * Twister + inversion, reached through unif_rand()/norm_rand(), P/inst/CppCode/Utils.cpp:61,
 * P/inst/CppCode/dists.cpp:954,1155) is NOT restated; both sides of every parity test consume this
 * stream instead, so that uniforms and normals are bit-identical on CPU and GPU ("replayed draws").
 *
 * Every floating-point step below is a single IEEE-754 operation (+ - * / sqrt fma); compile with
 * -ffp-contract=off so the compiler adds no fusion of its own.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include "draws_ref.h"

static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  int r;
  for (r = 0; r < 10; r++) {
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

static double bits2unif(uint32_t lo, uint32_t hi) {
  uint64_t m = (((uint64_t)hi << 32) | lo) >> 12; /* 52 bits */
  return ((double)m + 0.5) * 2.220446049250313e-16; /* 2^-52: (m+0.5) is exact, result in (0,1) */
}

void ref_unif2(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, uint32_t k,
               double *u1, double *u2) {
  uint32_t c[4];
  c[0] = (uint32_t)iter;
  c[1] = (sampler & 0xffffu) | ((uint32_t)((iter >> 32) & 0xffffu) << 16);
  c[2] = target;
  c[3] = k;
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32) ^ chain);
  *u1 = bits2unif(c[0], c[1]);
  *u2 = bits2unif(c[2], c[3]);
}

static double as_double(uint64_t b) { double d; memcpy(&d, &b, 8); return d; }
static uint64_t as_u64(double d) { uint64_t b; memcpy(&b, &d, 8); return b; }

double ref_det_log(double x) {
  uint64_t b = as_u64(x);
  int e = (int)(b >> 52) - 1023;
  double m = as_double((b & 0xFFFFFFFFFFFFFull) | 0x3FF0000000000000ull);
  double f, s, z, p, t;
  if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
  f = m - 1.0;
  s = f / (m + 1.0);
  z = s * s;
  p = 1.0 / 23.0;
  p = fma(p, z, 1.0 / 21.0);
  p = fma(p, z, 1.0 / 19.0);
  p = fma(p, z, 1.0 / 17.0);
  p = fma(p, z, 1.0 / 15.0);
  p = fma(p, z, 1.0 / 13.0);
  p = fma(p, z, 1.0 / 11.0);
  p = fma(p, z, 1.0 / 9.0);
  p = fma(p, z, 1.0 / 7.0);
  p = fma(p, z, 1.0 / 5.0);
  p = fma(p, z, 1.0 / 3.0);
  p = fma(p, z, 1.0);
  t = s * p;
  return fma((double)e, 0.6931471805599453, 2.0 * t);
}

double ref_det_exp(double x) {
  double kd, r, p;
  if (x < -700.0) return 0.0;
  kd = nearbyint(x * 1.4426950408889634);
  r = fma(-kd, 6.93147180369123816490e-01, x);
  r = fma(-kd, 1.90821492927058770002e-10, r);
  p = 1.0 / 87178291200.0;
  p = fma(p, r, 1.0 / 6227020800.0);
  p = fma(p, r, 1.0 / 479001600.0);
  p = fma(p, r, 1.0 / 39916800.0);
  p = fma(p, r, 1.0 / 3628800.0);
  p = fma(p, r, 1.0 / 362880.0);
  p = fma(p, r, 1.0 / 40320.0);
  p = fma(p, r, 1.0 / 5040.0);
  p = fma(p, r, 1.0 / 720.0);
  p = fma(p, r, 1.0 / 120.0);
  p = fma(p, r, 1.0 / 24.0);
  p = fma(p, r, 1.0 / 6.0);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  return p * as_double((uint64_t)((int64_t)kd + 1023) << 52);
}

static void det_sincos2pi(double u, double *s, double *c) {
  double a = u * 8.0;
  int o = (int)a;
  double f = a - (double)o;
  double th = f * 0.7853981633974483;
  double z = th * th;
  double ps, pc, sn, cs, t;
  ps = 1.0 / 355687428096000.0;           /* 1/17! */
  ps = fma(ps, z, -1.0 / 1307674368000.0); /* 15! */
  ps = fma(ps, z, 1.0 / 6227020800.0);     /* 13! */
  ps = fma(ps, z, -1.0 / 39916800.0);      /* 11! */
  ps = fma(ps, z, 1.0 / 362880.0);         /* 9! */
  ps = fma(ps, z, -1.0 / 5040.0);
  ps = fma(ps, z, 1.0 / 120.0);
  ps = fma(ps, z, -1.0 / 6.0);
  ps = fma(ps, z, 1.0);
  sn = th * ps;
  pc = -1.0 / 6402373705728000.0;          /* 18! */
  pc = fma(pc, z, 1.0 / 20922789888000.0); /* 16! */
  pc = fma(pc, z, -1.0 / 87178291200.0);   /* 14! */
  pc = fma(pc, z, 1.0 / 479001600.0);      /* 12! */
  pc = fma(pc, z, -1.0 / 3628800.0);       /* 10! */
  pc = fma(pc, z, 1.0 / 40320.0);
  pc = fma(pc, z, -1.0 / 720.0);
  pc = fma(pc, z, 1.0 / 24.0);
  pc = fma(pc, z, -0.5);
  pc = fma(pc, z, 1.0);
  cs = pc;
  if (o & 1) {
    double s2 = (sn + cs) * 0.7071067811865476;
    double c2 = (cs - sn) * 0.7071067811865476;
    sn = s2; cs = c2;
  }
  switch (o >> 1) {
    case 0: break;
    case 1: t = sn; sn = cs; cs = -t; break;
    case 2: sn = -sn; cs = -cs; break;
    default: t = sn; sn = -cs; cs = t; break;
  }
  *s = sn; *c = cs;
}

void ref_normal2(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, uint32_t k,
                 double *z0, double *z1) {
  double u1, u2, r, s, c;
  ref_unif2(seed, chain, iter, sampler, target, k, &u1, &u2);
  r = sqrt(-2.0 * ref_det_log(u1));
  det_sincos2pi(u2, &s, &c);
  *z0 = r * c;
  *z1 = r * s;
}

/* Marsaglia-Tsang gamma(a, 1) on the counter stream: call k=0 is the boost uniform (a<1),
 * trip m uses call 1+2m for the normal and 2+2m for the uniform. */
double ref_rgamma(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double a) {
  return ref_rgamma_at(seed, chain, iter, sampler, target, 0, a);
}
/* beta(a, b) = X / (X + Y), X ~ gamma(a) on calls 0.., Y ~ gamma(b) on calls 256.. of the same counter */
double ref_rbeta(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double a, double b) {
  double x = ref_rgamma_at(seed, chain, iter, sampler, target, 0, a);
  double y = ref_rgamma_at(seed, chain, iter, sampler, target, 256, b);
  return x / (x + y);
}
double ref_rgamma_at(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, uint32_t base, double a) {
  double boost = 1.0, d, c, x, v, u, x2, dummy, lhs, rhs;
  int m;
  if (!(a > 0.0) || !isfinite(a)) return NAN;
  if (a < 1.0) {
    ref_unif2(seed, chain, iter, sampler, target, base, &u, &dummy);
    boost = ref_det_exp(ref_det_log(u) / a);
    a = a + 1.0;
  }
  d = a - 1.0 / 3.0;
  c = (1.0 / 3.0) / sqrt(d);
  v = 1.0;
  for (m = 0; m < 64; m++) {
    ref_normal2(seed, chain, iter, sampler, target, base + 1 + 2 * m, &x, &dummy);
    v = 1.0 + c * x;
    if (v <= 0.0) continue;
    v = v * v * v;
    ref_unif2(seed, chain, iter, sampler, target, base + 2 + 2 * m, &u, &dummy);
    x2 = x * x;
    if (u < 1.0 - (0.0331 * x2) * x2) break;
    lhs = ref_det_log(u);
    rhs = 0.5 * x2 + d * ((1.0 - v) + ref_det_log(v));
    if (lhs < rhs) break;
  }
  return (d * v) * boost;
}

/* Poisson / binomial by inversion from zero (same sequence of operations as draws_dev.cuh: rpois_inv, rbinom_inv) */
double ref_rpois(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double mu) {
  double u, dummy, p, cdf;
  int x = 0;
  if (!(mu >= 0.0) || mu > 500.0) return NAN;
  if (mu == 0.0) return 0.0;
  ref_unif2(seed, chain, iter, sampler, target, 0, &u, &dummy);
  p = ref_det_exp(-mu); cdf = p;
  while (u > cdf && x < 5000) { x++; p = (p * mu) / (double)x; cdf = cdf + p; }
  return (double)x;
}
double ref_rbinom(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double n, double pr) {
  double u, dummy, q, r, p, cdf;
  int x = 0, ni;
  if (!(n >= 0.0) || n > 100000.0 || n != floor(n) || !(pr >= 0.0) || !(pr <= 1.0)) return NAN;
  if (pr == 0.0 || n == 0.0) return 0.0;
  if (pr == 1.0) return n;
  ref_unif2(seed, chain, iter, sampler, target, 0, &u, &dummy);
  q = 1.0 - pr; r = pr / q;
  p = ref_det_exp(n * ref_det_log(q));
  if (!(p > 0.0)) return NAN;
  cdf = p; ni = (int)n;
  while (u > cdf && x < ni) { p = ((p * (double)(ni - x)) / (double)(x + 1)) * r; x++; cdf = cdf + p; }
  return (double)x;
}

// nmath_dev.cuh -- device log-densities for the supported distribution set (sm_100a).
//
// The generated nodeFunction::calculate bodies of the reference call R's nmath d<dist>(x, ..., log=1)
// (call construction: P/R/nimbleFunction_nodeFunction.R:145-154; Rmath.h via P/inst/include/nimble/Utils.h:26-28).
// These are the same published algorithms (Loader's saddle-point dbinom/dpois: stirlerr + bd0) written for the
// device, FP64 throughout.  The translation unit that includes this header is compiled with -fmad=false so that
// every a*b+c stays two IEEE operations, as on the reference's x86-64 build (no FMA contraction at -O2).
#pragma once
#include <cfloat>
#include <cmath>

namespace nimb {

#define NIMB_LN_SQRT_2PI 0.918938533204672741780329736406
#define NIMB_LN_2PI 1.837877066409345483560659472811
#define NIMB_2PI 6.283185307179586476925286766559

// stirlerr(n/2), n = 0..30 (same values as nmath's sferr_halves; correctly rounded)
__device__ __constant__ double c_sferr_halves[31] = {
#include "sferr_table.inc"
};

__device__ __forceinline__ bool r_nonint(double x) { return fabs(x - nearbyint(x)) > 1e-7 * fmax(1.0, fabs(x)); }

__device__ __forceinline__ double stirlerr(double n) {
  const double S0 = 0.083333333333333333333, S1 = 0.00277777777777777777778, S2 = 0.00079365079365079365079365,
               S3 = 0.000595238095238095238095238, S4 = 0.0008417508417508417508417508;
  if (n <= 15.0) {
    double nn = n + n;
    if (nn == (double)(int)nn) return c_sferr_halves[(int)nn];
    return lgamma(n + 1.) - (n + 0.5) * log(n) + n - NIMB_LN_SQRT_2PI;
  }
  double nn = n * n;
  if (n > 500) return (S0 - S1 / nn) / n;
  if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
  if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
  return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

__device__ __forceinline__ double bd0(double x, double np) {
  if (!isfinite(x) || !isfinite(np) || np == 0.0) return NAN;
  if (fabs(x - np) < 0.1 * (x + np)) {
    double v = (x - np) / (x + np);
    double s = (x - np) * v;
    if (fabs(s) < DBL_MIN) return s;
    double ej = 2 * x * v;
    v = v * v;
    for (int j = 1; j < 1000; j++) {
      ej *= v;
      double s1 = s + ej / (double)((j << 1) + 1);
      if (s1 == s) return s1;
      s = s1;
    }
  }
  return x * log(x / np) + np - x;
}

__device__ __forceinline__ double dpois_raw(double x, double lambda) {
  if (lambda == 0) return (x == 0) ? 0. : -INFINITY;
  if (!isfinite(lambda)) return -INFINITY;
  if (x < 0) return -INFINITY;
  if (x <= lambda * DBL_MIN) return -lambda;
  if (lambda < x * DBL_MIN) return -lambda + x * log(lambda) - lgamma(x + 1);
  return -0.5 * log(NIMB_2PI * x) + (-stirlerr(x) - bd0(x, lambda));
}

__device__ __forceinline__ double dpois(double x, double lambda) {
  if (isnan(x) || isnan(lambda)) return x + lambda;
  if (lambda < 0) return NAN;
  if (r_nonint(x)) return -INFINITY;
  if (x < 0 || !isfinite(x)) return -INFINITY;
  return dpois_raw(nearbyint(x), lambda);
}

__device__ __forceinline__ double dbinom_raw(double x, double n, double p, double q) {
  if (p == 0) return (x == 0) ? 0. : -INFINITY;
  if (q == 0) return (x == n) ? 0. : -INFINITY;
  if (x == 0) {
    if (n == 0) return 0.;
    return (p < 0.1) ? -bd0(n, n * q) - n * p : n * log(q);
  }
  if (x == n) return (q < 0.1) ? -bd0(n, n * p) - n * q : n * log(p);
  if (x < 0 || x > n) return -INFINITY;
  double lc = stirlerr(n) - stirlerr(x) - stirlerr(n - x) - bd0(x, n * p) - bd0(n - x, n * q);
  double lf = NIMB_LN_2PI + log(x) + log1p(-x / n);
  return lc - 0.5 * lf;
}

__device__ __forceinline__ double dbinom(double x, double n, double p) {
  if (isnan(x) || isnan(n) || isnan(p)) return x + n + p;
  if (p < 0 || p > 1 || n < 0 || r_nonint(n)) return NAN;
  if (r_nonint(x)) return -INFINITY;
  if (x < 0 || !isfinite(x)) return -INFINITY;
  return dbinom_raw(nearbyint(x), nearbyint(n), p, 1 - p);
}

__device__ __forceinline__ double dnorm(double x, double mu, double sigma) {
  if (isnan(x) || isnan(mu) || isnan(sigma)) return x + mu + sigma;
  if (sigma < 0) return NAN;
  if (!isfinite(sigma)) return -INFINITY;
  if (!isfinite(x) && mu == x) return NAN;
  if (sigma == 0) return (x == mu) ? INFINITY : -INFINITY;
  x = (x - mu) / sigma;
  if (!isfinite(x)) return -INFINITY;
  x = fabs(x);
  if (x >= 2 * sqrt(DBL_MAX)) return -INFINITY;
  return -(NIMB_LN_SQRT_2PI + 0.5 * x * x + log(sigma));
}

__device__ __forceinline__ double dgamma(double x, double shape, double scale) {
  if (isnan(x) || isnan(shape) || isnan(scale)) return x + shape + scale;
  if (shape < 0 || scale <= 0) return NAN;
  if (x < 0) return -INFINITY;
  if (shape == 0) return (x == 0) ? INFINITY : -INFINITY;
  if (x == 0) {
    if (shape < 1) return INFINITY;
    if (shape > 1) return -INFINITY;
    return -log(scale);
  }
  if (shape < 1) {
    double pr = dpois_raw(shape, x / scale);
    return pr + (isfinite(shape / x) ? log(shape / x) : log(shape) - log(x));
  }
  return dpois_raw(shape - 1, x / scale) - log(scale);
}

__device__ __forceinline__ double lgammacor(double x) {
  double x2 = x * x;
  return (1. / 12 - (1. / 360 - (1. / 1260 - (1. / 1680 - (1. / 1188 - (691. / 360360 - (1. / 156) / x2) / x2) / x2) / x2) / x2) / x2) / x;
}

__device__ __forceinline__ double lbeta(double a, double b) {
  double p = fmin(a, b), q = fmax(a, b);
  if (p < 0) return NAN;
  if (p == 0) return INFINITY;
  if (!isfinite(q)) return -INFINITY;
  if (p >= 10) {
    double corr = lgammacor(p) + lgammacor(q) - lgammacor(p + q);
    return log(q) * -0.5 + NIMB_LN_SQRT_2PI + corr + (p - 0.5) * log(p / (p + q)) + q * log1p(-p / (p + q));
  } else if (q >= 10) {
    double corr = lgammacor(q) - lgammacor(p + q);
    return lgamma(p) + corr + p - p * log(p + q) + (q - 0.5) * log1p(-p / (p + q));
  }
  return lgamma(p) + lgamma(q) - lgamma(p + q);
}

__device__ __forceinline__ double dbeta(double x, double a, double b) {
  if (isnan(x) || isnan(a) || isnan(b)) return x + a + b;
  if (a < 0 || b < 0) return NAN;
  if (x < 0 || x > 1) return -INFINITY;
  if (a == 0 || b == 0 || !isfinite(a) || !isfinite(b)) {
    if (a == 0 && b == 0) return (x == 0 || x == 1) ? INFINITY : -INFINITY;
    if (a == 0 || a / b == 0) return (x == 0) ? INFINITY : -INFINITY;
    if (b == 0 || b / a == 0) return (x == 1) ? INFINITY : -INFINITY;
    return (x == 0.5) ? INFINITY : -INFINITY;
  }
  if (x == 0) {
    if (a > 1) return -INFINITY;
    if (a < 1) return INFINITY;
    return log(b);
  }
  if (x == 1) {
    if (b > 1) return -INFINITY;
    if (b < 1) return INFINITY;
    return log(a);
  }
  if (a <= 2 || b <= 2) return (a - 1) * log(x) + (b - 1) * log1p(-x) - lbeta(a, b);
  return log(a + b - 1) + dbinom_raw(a - 1, a + b - 2, x, 1 - x);
}

__device__ __forceinline__ double dlogis(double x, double location, double scale) {
  if (isnan(x) || isnan(location) || isnan(scale)) return x + location + scale;
  if (scale <= 0.0) return NAN;
  x = fabs((x - location) / scale);
  double e = exp(-x);
  double f = 1.0 + e;
  return -(x + log(scale * f * f));
}

// [nmath/dlnorm.c]
__device__ __forceinline__ double dlnorm(double x, double meanlog, double sdlog) {
  if (isnan(x) || isnan(meanlog) || isnan(sdlog)) return x + meanlog + sdlog;
  if (sdlog < 0) return NAN;
  if (!isfinite(x) && log(x) == meanlog) return NAN;
  if (sdlog == 0) return (log(x) == meanlog) ? INFINITY : -INFINITY;
  if (x <= 0) return -INFINITY;
  double y = (log(x) - meanlog) / sdlog;
  return -(NIMB_LN_SQRT_2PI + 0.5 * y * y + log(x * sdlog));
}

// dexp_nimble(x, rate) = dexp(x, 1/rate): P/inst/CppCode/dists.cpp:1804-1807
__device__ __forceinline__ double dexp_nimble(double x, double rate) {
  double scale = 1 / rate;
  if (isnan(x) || isnan(scale)) return x + scale;
  if (scale <= 0.0) return NAN;
  if (x < 0.) return -INFINITY;
  return (-x / scale) - log(scale);
}

}  // namespace nimb

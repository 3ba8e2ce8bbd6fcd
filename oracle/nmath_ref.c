/*
 * nmath_ref.c -- TEST INFRASTRUCTURE (parity oracle).  Not part of the product; nothing under
 * nimble_b200/ may link or call it.
 *
 * CPU restatement of the scalar log-densities that NIMBLE's generated nodeFunction::calculate bodies
 * call (reference call sites: P/R/nimbleFunction_nodeFunction.R:145-154 build `d<dist>(x, params, 1)`
 * from distInfo$density_code; Rmath.h is pulled in at P/inst/include/nimble/Utils.h:26-28).
 *
 * The arithmetic itself lives in a third-party dependency that is NOT under /root/reference:
 * R's nmath (libR / Rmath), pinned by the reference only as `Depends: R (>= 3.1.2)` (P/DESCRIPTION).
 * This file restates the published nmath algorithms of that era (R 3.1.2 .. 4.0.x: Catherine Loader's
 * saddle-point dbinom/dpois with stirlerr + bd0): dnorm4, dpois_raw, dbinom_raw, dgamma, dbeta, dlogis.
 * Deviations, both documented in DESIGN.md: (1) lgammafn/gammafn use the platform lgamma() instead of
 * R's Chebyshev kernels (coefficients not reproducible without the source; <= 2 ulp), (2) lgammacor is
 * evaluated from the Stirling series.  dcat follows P/inst/CppCode/dists.cpp:914-934, which IS in tree.
 *
 * PARITY PINNING: the reference holds no literal golden values for these densities (its tests compare
 * to the running R's own nmath, P/tests/testthat/test-dsl_dists.R:61-66); they are pinned here against
 * 50-digit mpmath evaluations (tests/golden/density_vectors.json, made by tests/golden/make_density_vectors.py).
 */
#include <float.h>
#include <math.h>
#include "nmath_ref.h"
#include "stirlerr_table.h"

#define M_LN_SQRT_2PI_ 0.918938533204672741780329736406 /* log(sqrt(2*pi)) */
#define M_LN_2PI_ 1.837877066409345483560659472811      /* log(2*pi) */
#define M_2PI_ 6.283185307179586476925286766559
#define ML_NEGINF (-INFINITY)
#define ML_POSINF (INFINITY)

static int r_nonint(double x) { return fabs(x - nearbyint(x)) > 1e-7 * fmax(1.0, fabs(x)); }

double ref_lgammafn(double x) { return lgamma(x); }

/* stirlerr(n) = log(n!) - log( sqrt(2*pi*n)*(n/e)^n )   [nmath/stirlerr.c] */
double ref_stirlerr(double n) {
  const double S0 = 0.083333333333333333333;         /* 1/12 */
  const double S1 = 0.00277777777777777777778;       /* 1/360 */
  const double S2 = 0.00079365079365079365079365;    /* 1/1260 */
  const double S3 = 0.000595238095238095238095238;   /* 1/1680 */
  const double S4 = 0.0008417508417508417508417508;  /* 1/1188 */
  double nn;
  if (n <= 15.0) {
    nn = n + n;
    if (nn == (int)nn) return sferr_halves[(int)nn];
    return ref_lgammafn(n + 1.) - (n + 0.5) * log(n) + n - M_LN_SQRT_2PI_;
  }
  nn = n * n;
  if (n > 500) return (S0 - S1 / nn) / n;
  if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
  if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
  return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

/* bd0(x, np) = x log(x/np) + np - x, series for x ~ np   [nmath/bd0.c] */
double ref_bd0(double x, double np) {
  double ej, s, s1, v;
  int j;
  if (!isfinite(x) || !isfinite(np) || np == 0.0) return NAN;
  if (fabs(x - np) < 0.1 * (x + np)) {
    v = (x - np) / (x + np);
    s = (x - np) * v;
    if (fabs(s) < DBL_MIN) return s;
    ej = 2 * x * v;
    v = v * v;
    for (j = 1; j < 1000; j++) {
      ej *= v;
      s1 = s + ej / ((j << 1) + 1);
      if (s1 == s) return s1;
      s = s1;
    }
  }
  return x * log(x / np) + np - x;
}

/* [nmath/dpois.c] */
double ref_dpois_raw(double x, double lambda) {
  if (lambda == 0) return (x == 0) ? 0. : ML_NEGINF;
  if (!isfinite(lambda)) return ML_NEGINF;
  if (x < 0) return ML_NEGINF;
  if (x <= lambda * DBL_MIN) return -lambda;
  if (lambda < x * DBL_MIN) return -lambda + x * log(lambda) - ref_lgammafn(x + 1);
  return -0.5 * log(M_2PI_ * x) + (-ref_stirlerr(x) - ref_bd0(x, lambda));
}

double ref_dpois(double x, double lambda) {
  if (isnan(x) || isnan(lambda)) return x + lambda;
  if (lambda < 0) return NAN;
  if (r_nonint(x)) return ML_NEGINF;
  if (x < 0 || !isfinite(x)) return ML_NEGINF;
  x = nearbyint(x);
  return ref_dpois_raw(x, lambda);
}

/* [nmath/dbinom.c] */
double ref_dbinom_raw(double x, double n, double p, double q) {
  double lf, lc;
  if (p == 0) return (x == 0) ? 0. : ML_NEGINF;
  if (q == 0) return (x == n) ? 0. : ML_NEGINF;
  if (x == 0) {
    if (n == 0) return 0.;
    lc = (p < 0.1) ? -ref_bd0(n, n * q) - n * p : n * log(q);
    return lc;
  }
  if (x == n) {
    lc = (q < 0.1) ? -ref_bd0(n, n * p) - n * q : n * log(p);
    return lc;
  }
  if (x < 0 || x > n) return ML_NEGINF;
  lc = ref_stirlerr(n) - ref_stirlerr(x) - ref_stirlerr(n - x) - ref_bd0(x, n * p) - ref_bd0(n - x, n * q);
  lf = M_LN_2PI_ + log(x) + log1p(-x / n);
  return lc - 0.5 * lf;
}

double ref_dbinom(double x, double n, double p) {
  if (isnan(x) || isnan(n) || isnan(p)) return x + n + p;
  if (p < 0 || p > 1 || n < 0 || r_nonint(n)) return NAN;
  if (r_nonint(x)) return ML_NEGINF;
  if (x < 0 || !isfinite(x)) return ML_NEGINF;
  n = nearbyint(n);
  x = nearbyint(x);
  return ref_dbinom_raw(x, n, p, 1 - p);
}

/* [nmath/dnorm.c] dnorm4 */
double ref_dnorm(double x, double mu, double sigma) {
  if (isnan(x) || isnan(mu) || isnan(sigma)) return x + mu + sigma;
  if (sigma < 0) return NAN;
  if (!isfinite(sigma)) return ML_NEGINF;
  if (!isfinite(x) && mu == x) return NAN;
  if (sigma == 0) return (x == mu) ? ML_POSINF : ML_NEGINF;
  x = (x - mu) / sigma;
  if (!isfinite(x)) return ML_NEGINF;
  x = fabs(x);
  if (x >= 2 * sqrt(DBL_MAX)) return ML_NEGINF;
  return -(M_LN_SQRT_2PI_ + 0.5 * x * x + log(sigma));
}

/* [nmath/dgamma.c] */
double ref_dgamma(double x, double shape, double scale) {
  double pr;
  if (isnan(x) || isnan(shape) || isnan(scale)) return x + shape + scale;
  if (shape < 0 || scale <= 0) return NAN;
  if (x < 0) return ML_NEGINF;
  if (shape == 0) return (x == 0) ? ML_POSINF : ML_NEGINF;
  if (x == 0) {
    if (shape < 1) return ML_POSINF;
    if (shape > 1) return ML_NEGINF;
    return -log(scale);
  }
  if (shape < 1) {
    pr = ref_dpois_raw(shape, x / scale);
    return pr + (isfinite(shape / x) ? log(shape / x) : log(shape) - log(x));
  }
  pr = ref_dpois_raw(shape - 1, x / scale);
  return pr - log(scale);
}

/* lgammacor(x) = lgamma(x) - (log(sqrt(2pi)) + (x-.5)log(x) - x), x >= 10: Stirling series */
static double ref_lgammacor(double x) {
  double x2 = x * x;
  return (1. / 12 - (1. / 360 - (1. / 1260 - (1. / 1680 - (1. / 1188 - (691. / 360360 - (1. / 156) / x2) / x2) / x2) / x2) / x2) / x2) / x;
}

/* [nmath/lbeta.c] */
double ref_lbeta(double a, double b) {
  double corr, p, q;
  p = q = a;
  if (b < p) p = b;
  if (b > q) q = b;
  if (p < 0) return NAN;
  if (p == 0) return ML_POSINF;
  if (!isfinite(q)) return ML_NEGINF;
  if (p >= 10) {
    corr = ref_lgammacor(p) + ref_lgammacor(q) - ref_lgammacor(p + q);
    return log(q) * -0.5 + M_LN_SQRT_2PI_ + corr + (p - 0.5) * log(p / (p + q)) + q * log1p(-p / (p + q));
  } else if (q >= 10) {
    corr = ref_lgammacor(q) - ref_lgammacor(p + q);
    return ref_lgammafn(p) + corr + p - p * log(p + q) + (q - 0.5) * log1p(-p / (p + q));
  }
  return ref_lgammafn(p) + ref_lgammafn(q) - ref_lgammafn(p + q);
}

/* [nmath/dbeta.c] */
double ref_dbeta(double x, double a, double b) {
  if (isnan(x) || isnan(a) || isnan(b)) return x + a + b;
  if (a < 0 || b < 0) return NAN;
  if (x < 0 || x > 1) return ML_NEGINF;
  if (a == 0 || b == 0 || !isfinite(a) || !isfinite(b)) {
    if (a == 0 && b == 0) return (x == 0 || x == 1) ? ML_POSINF : ML_NEGINF;
    if (a == 0 || a / b == 0) return (x == 0) ? ML_POSINF : ML_NEGINF;
    if (b == 0 || b / a == 0) return (x == 1) ? ML_POSINF : ML_NEGINF;
    return (x == 0.5) ? ML_POSINF : ML_NEGINF;
  }
  if (x == 0) {
    if (a > 1) return ML_NEGINF;
    if (a < 1) return ML_POSINF;
    return log(b);
  }
  if (x == 1) {
    if (b > 1) return ML_NEGINF;
    if (b < 1) return ML_POSINF;
    return log(a);
  }
  if (a <= 2 || b <= 2) return (a - 1) * log(x) + (b - 1) * log1p(-x) - ref_lbeta(a, b);
  return log(a + b - 1) + ref_dbinom_raw(a - 1, a + b - 2, x, 1 - x);
}

/* [nmath/dlogis.c] */
double ref_dlogis(double x, double location, double scale) {
  double e, f;
  if (isnan(x) || isnan(location) || isnan(scale)) return x + location + scale;
  if (scale <= 0.0) return NAN;
  x = fabs((x - location) / scale);
  e = exp(-x);
  f = 1.0 + e;
  return -(x + log(scale * f * f));
}

/* [nmath/dlnorm.c] */
double ref_dlnorm(double x, double meanlog, double sdlog) {
  double y;
  if (isnan(x) || isnan(meanlog) || isnan(sdlog)) return x + meanlog + sdlog;
  if (sdlog < 0) return NAN;
  if (!isfinite(x) && log(x) == meanlog) return NAN; /* log(x) - meanlog is NaN */
  if (sdlog == 0) return (log(x) == meanlog) ? ML_POSINF : ML_NEGINF;
  if (x <= 0) return ML_NEGINF;
  y = (log(x) - meanlog) / sdlog;
  return -(M_LN_SQRT_2PI_ + 0.5 * y * y + log(x * sdlog));
}

/* dexp_nimble(x, rate) = dexp(x, 1/rate): P/inst/CppCode/dists.cpp:1804-1807; [nmath/dexp.c] */
double ref_dexp_nimble(double x, double rate) {
  double scale = 1 / rate;
  if (isnan(x) || isnan(scale)) return x + scale;
  if (scale <= 0.0) return NAN;
  if (x < 0.) return ML_NEGINF;
  return (-x / scale) - log(scale);
}

/* P/inst/CppCode/dists.cpp:914-934 */
double ref_dcat(double x, const double *prob, int K) {
  int i;
  double sumProb = 0.0;
  if (isnan(x)) return x;
  for (i = 0; i < K; i++)
    if (isnan(prob[i])) return prob[i];
  for (i = 0; i < K; i++) {
    if (prob[i] < 0) return NAN;
    sumProb += prob[i];
  }
  if (r_nonint(x)) return ML_NEGINF;
  x = nearbyint(x);
  if (x > K || x < 1) return ML_NEGINF;
  return log(prob[(int)x - 1]) - log(sumProb);
}

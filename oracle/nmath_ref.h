/* nmath_ref.h -- TEST INFRASTRUCTURE (parity oracle); see nmath_ref.c */
#ifndef NMATH_REF_H
#define NMATH_REF_H
#ifdef __cplusplus
extern "C" {
#endif
double ref_lgammafn(double x);
double ref_stirlerr(double n);
double ref_bd0(double x, double np);
double ref_dpois_raw(double x, double lambda);
double ref_dpois(double x, double lambda);
double ref_dbinom_raw(double x, double n, double p, double q);
double ref_dbinom(double x, double n, double p);
double ref_dnorm(double x, double mu, double sigma);
double ref_dgamma(double x, double shape, double scale);
double ref_lbeta(double a, double b);
double ref_dbeta(double x, double a, double b);
double ref_dlogis(double x, double location, double scale);
double ref_dexp_nimble(double x, double rate);
double ref_dlnorm(double x, double meanlog, double sdlog);
double ref_dcat(double x, const double *prob, int K);
#ifdef __cplusplus
}
#endif
#endif

/* draws_ref.h -- TEST INFRASTRUCTURE (parity oracle); see draws_ref.c */
#ifndef DRAWS_REF_H
#define DRAWS_REF_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
void ref_unif2(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, uint32_t k,
               double *u1, double *u2);
void ref_normal2(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, uint32_t k,
                 double *z0, double *z1);
double ref_rgamma(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double a);
double ref_rgamma_at(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, uint32_t base, double a);
double ref_rbeta(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double a, double b);
double ref_rpois(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double mu);
double ref_rbinom(uint64_t seed, uint32_t chain, uint64_t iter, uint32_t sampler, uint32_t target, double n, double pr);
double ref_det_log(double x);
double ref_det_exp(double x);
#ifdef __cplusplus
}
#endif
#endif

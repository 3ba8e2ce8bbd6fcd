/* nimb200_R.c -- the .Call shim between R and libnimb200.so (include/nimb200.h).
 *
 * UNVERIFIED: neither R nor Rinternals.h exists in the build image, so this file has never been compiled.  It follows the
 * reference's own convention for compiled entry points -- extern "C" SEXP CALL_<name>(SEXP args..., SEXP SextPtrToObject),
 * the object recovered with R_ExternalPtrAddr, a finalizer registered on the external pointer, errors through Rf_error
 * (P/R/cppDefs_RCfunction.R:206-332; P/inst/CppCode/dllFinalizer.cpp; NIMERROR = Rf_error, P/inst/include/nimble/Utils.h:60).
 * The Python binding nimble_b200/_capi.py drives exactly the same C entry points and is what the tests run.
 *
 * Build (once R is there):  R CMD SHLIB nimb200_R.c -I../../include -L../../nimble_b200 -lnimb200
 *
 * The node table arrives as the R list lower_model() builds (r/R/lower_model.R):
 *   n_slots, darr (list of double vectors), iarr (list of integer vectors),
 *   ops (integer matrix [4 x n_ops]: op, a, b, c) + ops_k (double[n_ops]),
 *   decl (integer matrix [18 x n_decl], rows in the order of struct nimb200_decl),
 *   samplers_i (integer matrix [14 x n_samplers]) + samplers_d (double matrix [4 x n_samplers]),
 *   deps (integer matrix [8 x n_deps]), monitor_slots (integer) -- all indices 0-based already. */
#include <R.h>
#include <Rinternals.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "nimb200.h"

typedef struct {
  nimb200_table t;
  const double **darr; int64_t *darr_len;
  const int32_t **iarr; int64_t *iarr_len;
  nimb200_op *ops; nimb200_decl *decl; nimb200_sampler *samplers; nimb200_dep *deps;
} table_box;

static SEXP list_get(SEXP list, const char *name) {
  SEXP names = getAttrib(list, R_NamesSymbol);
  for (R_xlen_t i = 0; i < XLENGTH(list); i++)
    if (strcmp(CHAR(STRING_ELT(names, i)), name) == 0) return VECTOR_ELT(list, i);
  Rf_error("node table: element '%s' is missing", name);
  return R_NilValue;
}

static void free_box(table_box *b) {
  free((void *)b->darr); free(b->darr_len); free((void *)b->iarr); free(b->iarr_len);
  free(b->ops); free(b->decl); free(b->samplers); free(b->deps);
}

/* Copies the small structs; the big arrays (data, constants, index vectors) are NOT copied: the table points at REAL() /
 * INTEGER() of the list elements, which stay protected by the caller for the duration of nimb200_model_create (the library
 * uploads what it keeps and never holds on to host pointers). */
static void fill_table(SEXP tab, table_box *b) {
  memset(b, 0, sizeof *b);
  SEXP darr = list_get(tab, "darr"), iarr = list_get(tab, "iarr");
  SEXP ops = list_get(tab, "ops"), ops_k = list_get(tab, "ops_k"), decl = list_get(tab, "decl");
  SEXP si = list_get(tab, "samplers_i"), sd = list_get(tab, "samplers_d"), deps = list_get(tab, "deps");
  SEXP mon = list_get(tab, "monitor_slots");
  nimb200_table *t = &b->t;
  t->version = NIMB200_TABLE_VERSION;
  t->n_slots = asInteger(list_get(tab, "n_slots"));
  t->n_darr = (int32_t)XLENGTH(darr);
  t->n_iarr = (int32_t)XLENGTH(iarr);
  b->darr = (const double **)calloc(t->n_darr + 1, sizeof(double *));
  b->darr_len = (int64_t *)calloc(t->n_darr + 1, sizeof(int64_t));
  b->iarr = (const int32_t **)calloc(t->n_iarr + 1, sizeof(int32_t *));
  b->iarr_len = (int64_t *)calloc(t->n_iarr + 1, sizeof(int64_t));
  for (int i = 0; i < t->n_darr; i++) { b->darr[i] = REAL(VECTOR_ELT(darr, i)); b->darr_len[i] = XLENGTH(VECTOR_ELT(darr, i)); }
  for (int i = 0; i < t->n_iarr; i++) { b->iarr[i] = (const int32_t *)INTEGER(VECTOR_ELT(iarr, i)); b->iarr_len[i] = XLENGTH(VECTOR_ELT(iarr, i)); }
  t->darr = b->darr; t->darr_len = b->darr_len; t->iarr = b->iarr; t->iarr_len = b->iarr_len;
  t->n_ops = (int32_t)XLENGTH(ops_k);
  b->ops = (nimb200_op *)calloc(t->n_ops + 1, sizeof(nimb200_op));
  for (int i = 0; i < t->n_ops; i++) {
    const int *o = INTEGER(ops) + 4 * i;
    b->ops[i].op = o[0]; b->ops[i].a = o[1]; b->ops[i].b = o[2]; b->ops[i].c = o[3]; b->ops[i].k = REAL(ops_k)[i];
  }
  t->ops = b->ops;
  t->n_decl = (int32_t)(XLENGTH(decl) / 18);
  b->decl = (nimb200_decl *)calloc(t->n_decl + 1, sizeof(nimb200_decl));
  for (int i = 0; i < t->n_decl; i++) memcpy(&b->decl[i], INTEGER(decl) + 18 * i, 18 * sizeof(int32_t));   /* 18 int32 fields, same order */
  t->decl = b->decl;
  t->n_samplers = (int32_t)(XLENGTH(si) / 14);
  b->samplers = (nimb200_sampler *)calloc(t->n_samplers + 1, sizeof(nimb200_sampler));
  for (int i = 0; i < t->n_samplers; i++) {
    memcpy(&b->samplers[i], INTEGER(si) + 14 * i, 14 * sizeof(int32_t));
    const double *d = REAL(sd) + 4 * i;
    b->samplers[i].scale = d[0]; b->samplers[i].adapt_factor_exponent = d[1]; b->samplers[i].lower = d[2]; b->samplers[i].upper = d[3];
  }
  t->samplers = b->samplers;
  t->n_deps = (int32_t)(XLENGTH(deps) / 8);
  b->deps = (nimb200_dep *)calloc(t->n_deps + 1, sizeof(nimb200_dep));
  for (int i = 0; i < t->n_deps; i++) memcpy(&b->deps[i], INTEGER(deps) + 8 * i, 8 * sizeof(int32_t));
  t->deps = b->deps;
  t->n_monitors = (int32_t)XLENGTH(mon);
  t->monitor_slots = (const int32_t *)INTEGER(mon);
}

static void model_finalizer(SEXP ptr) {
  nimb200_model *m = (nimb200_model *)R_ExternalPtrAddr(ptr);
  if (m) { nimb200_model_free(m); R_ClearExternalPtr(ptr); }
}

static nimb200_model *model_of(SEXP ptr) {
  nimb200_model *m = (nimb200_model *)R_ExternalPtrAddr(ptr);
  if (!m) Rf_error("this MCMC has been cleared");            /* cppDefs_RCfunction.R:171-177 */
  return m;
}

/* compileNimble(mcmc) for the device back end: table -> external pointer that owns every device buffer */
SEXP CALL_nimb200_build(SEXP table, SEXP nchains, SEXP first_chain, SEXP device) {
  table_box box;
  nimb200_model *m = NULL;
  fill_table(table, &box);
  int rc = nimb200_model_create(&box.t, asInteger(nchains), asInteger(first_chain), asInteger(device), &m);
  free_box(&box);                                             /* before any longjmp */
  if (rc) Rf_error("%s", nimb200_last_error());
  SEXP ptr = PROTECT(R_MakeExternalPtr(m, R_NilValue, R_NilValue));
  R_RegisterCFinalizerEx(ptr, model_finalizer, TRUE);
  UNPROTECT(1);
  return ptr;
}

/* Cmcmc$run for all chains at once.  inits: double matrix [n_slots x nchains] (R is column-major, so this IS the
 * [nchains][n_slots] host layout of nimb200_set_values) or NULL to continue from the resident state. */
SEXP CALL_nimb200_run(SEXP ptr, SEXP inits, SEXP decl_order, SEXP niter, SEXP nburnin, SEXP thin, SEXP reset, SEXP seed) {
  nimb200_model *m = model_of(ptr);
  uint64_t s = (uint64_t)asReal(seed);
  if (!isNull(inits)) {
    if (nimb200_set_values(m, REAL(inits))) Rf_error("%s", nimb200_last_error());
    /* initializeModel: NA initial values are simulated from their distributions, parents first (initializeModel.R:78-139) */
    int any_na = 0;
    for (R_xlen_t i = 0; i < XLENGTH(inits) && !any_na; i++) any_na = ISNAN(REAL(inits)[i]);
    if (any_na && nimb200_initialize_model(m, (const int32_t *)INTEGER(decl_order), (int32_t)XLENGTH(decl_order), s))
      Rf_error("%s", nimb200_last_error());
  }
  if (nimb200_mcmc_run(m, (int64_t)asReal(niter), (int64_t)asReal(nburnin), asInteger(thin), asLogical(reset), s))
    Rf_error("%s", nimb200_last_error());
  return R_NilValue;
}

/* as.matrix(mvSamples) of every chain: array [col, row, chain]; aperm(x, c(2, 1, 3))[, , k] is chain k's matrix */
SEXP CALL_nimb200_samples(SEXP ptr, SEXP nmon, SEXP nchains) {
  nimb200_model *m = model_of(ptr);
  int64_t rows = 0;
  if (nimb200_mcmc_samples(m, NULL, &rows)) Rf_error("%s", nimb200_last_error());
  SEXP out = PROTECT(alloc3DArray(REALSXP, asInteger(nmon), (int)rows, asInteger(nchains)));
  if (rows > 0 && nimb200_mcmc_samples(m, REAL(out), &rows)) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  UNPROTECT(1);
  return out;
}

/* model$getValues / setValues for all chains: [n_slots x nchains] */
SEXP CALL_nimb200_get_state(SEXP ptr, SEXP n_slots, SEXP nchains) {
  nimb200_model *m = model_of(ptr);
  SEXP out = PROTECT(allocMatrix(REALSXP, asInteger(n_slots), asInteger(nchains)));
  if (nimb200_get_values(m, REAL(out))) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  UNPROTECT(1);
  return out;
}
SEXP CALL_nimb200_set_state(SEXP ptr, SEXP values) {
  if (nimb200_set_values(model_of(ptr), REAL(values))) Rf_error("%s", nimb200_last_error());
  return R_NilValue;
}

/* sampler member variables (scale, timesAdapted, propCov): what = 0, 1, 2; length = number of doubles expected */
SEXP CALL_nimb200_sampler_state(SEXP ptr, SEXP sampler, SEXP what, SEXP length) {
  nimb200_model *m = model_of(ptr);
  R_xlen_t n = (R_xlen_t)asReal(length);
  SEXP out = PROTECT(allocVector(REALSXP, n));
  if (nimb200_sampler_state(m, asInteger(sampler), asInteger(what), REAL(out), (int64_t)n)) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  UNPROTECT(1);
  return out;
}

/* run(time = TRUE) / getTimes() (MCMC_build.R:281-285, 323-326) */
SEXP CALL_nimb200_set_timing(SEXP ptr, SEXP enable) {
  if (nimb200_mcmc_set_timing(model_of(ptr), asLogical(enable))) Rf_error("%s", nimb200_last_error());
  return R_NilValue;
}
SEXP CALL_nimb200_times(SEXP ptr, SEXP n_samplers) {
  int n = asInteger(n_samplers);
  SEXP out = PROTECT(allocVector(REALSXP, n));
  if (nimb200_mcmc_sampler_times(model_of(ptr), REAL(out), n)) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  for (int i = 0; i < n; i++) REAL(out)[i] *= 1e-3;           /* getTimes() reports seconds */
  UNPROTECT(1);
  return out;
}

/* model$calculate(nodes) on one whole declaration, one value per chain */
SEXP CALL_nimb200_calculate(SEXP ptr, SEXP decl, SEXP n_nodes, SEXP nchains) {
  SEXP out = PROTECT(allocVector(REALSXP, asInteger(nchains)));
  if (nimb200_calculate(model_of(ptr), asInteger(decl), 0, asInteger(n_nodes), REAL(out))) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  UNPROTECT(1);
  return out;
}

/* online WAIC: configureMCMC(enableWAIC = TRUE) / mcmc$getWAIC() (MCMC_WAIC.R:209-300) */
SEXP CALL_nimb200_waic_enable(SEXP ptr, SEXP enable) {
  if (nimb200_waic_enable(model_of(ptr), asLogical(enable))) Rf_error("%s", nimb200_last_error());
  return R_NilValue;
}
SEXP CALL_nimb200_waic_get(SEXP ptr) {   /* c(WAIC, lppd, pWAIC, n) */
  SEXP out = PROTECT(allocVector(REALSXP, 4));
  if (nimb200_waic_get(model_of(ptr), REAL(out))) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  UNPROTECT(1);
  return out;
}

/* samplesSummary (MCMC_utils.R:315-341) computed on the device over chains [chain_lo, chain_hi): 5 x n_monitors matrix
 * (Mean, Median, St.Dev., 95%CI_low, 95%CI_upp), column-major as R stores it */
SEXP CALL_nimb200_samples_summary(SEXP ptr, SEXP chain_lo, SEXP chain_hi, SEXP nmon) {
  int n = asInteger(nmon);
  SEXP buf = PROTECT(allocVector(REALSXP, (R_xlen_t)5 * n));      /* the library writes [5][nmon], row-major */
  if (nimb200_samples_summary(model_of(ptr), asInteger(chain_lo), asInteger(chain_hi), REAL(buf))) { UNPROTECT(1); Rf_error("%s", nimb200_last_error()); }
  SEXP out = PROTECT(allocMatrix(REALSXP, 5, n));
  for (int i = 0; i < 5; i++)
    for (int j = 0; j < n; j++) REAL(out)[i + 5 * (R_xlen_t)j] = REAL(buf)[(R_xlen_t)i * n + j];
  UNPROTECT(2);
  return out;
}

/* nimbleOptions(MCMCsaveHistory = TRUE): getScaleHistory / getAcceptanceHistory / getPropCovHistory (MCMC_samplers.R:394-448,
 * 726-780).  what = 0, 1, 2; per_adaptation = doubles per recorded adaptation (chains, or d*d*chains for propCov) */
SEXP CALL_nimb200_history_enable(SEXP ptr, SEXP max_adaptations) {
  if (nimb200_sampler_history_enable(model_of(ptr), asInteger(max_adaptations))) Rf_error("%s", nimb200_last_error());
  return R_NilValue;
}
SEXP CALL_nimb200_history(SEXP ptr, SEXP sampler, SEXP what, SEXP per_adaptation) {
  nimb200_model *m = model_of(ptr);
  int32_t n_adapt = 0;
  if (nimb200_sampler_history(m, asInteger(sampler), asInteger(what), NULL, 0, &n_adapt)) Rf_error("%s", nimb200_last_error());
  R_xlen_t n = (R_xlen_t)n_adapt * (R_xlen_t)asReal(per_adaptation);
  SEXP out = PROTECT(allocVector(REALSXP, n));
  if (n > 0 && nimb200_sampler_history(m, asInteger(sampler), asInteger(what), REAL(out), (int64_t)n, &n_adapt)) {
    UNPROTECT(1);
    Rf_error("%s", nimb200_last_error());
  }
  UNPROTECT(1);
  return out;
}

/* explicit release (the finalizer does the same at garbage collection) */
SEXP CALL_nimb200_clear(SEXP ptr) {
  model_finalizer(ptr);
  return R_NilValue;
}

#include <R_ext/Rdynload.h>
static const R_CallMethodDef call_methods[] = {
  {"CALL_nimb200_build", (DL_FUNC)&CALL_nimb200_build, 4},
  {"CALL_nimb200_run", (DL_FUNC)&CALL_nimb200_run, 8},
  {"CALL_nimb200_samples", (DL_FUNC)&CALL_nimb200_samples, 3},
  {"CALL_nimb200_get_state", (DL_FUNC)&CALL_nimb200_get_state, 3},
  {"CALL_nimb200_set_state", (DL_FUNC)&CALL_nimb200_set_state, 2},
  {"CALL_nimb200_sampler_state", (DL_FUNC)&CALL_nimb200_sampler_state, 4},
  {"CALL_nimb200_set_timing", (DL_FUNC)&CALL_nimb200_set_timing, 2},
  {"CALL_nimb200_times", (DL_FUNC)&CALL_nimb200_times, 2},
  {"CALL_nimb200_calculate", (DL_FUNC)&CALL_nimb200_calculate, 4},
  {"CALL_nimb200_waic_enable", (DL_FUNC)&CALL_nimb200_waic_enable, 2},
  {"CALL_nimb200_waic_get", (DL_FUNC)&CALL_nimb200_waic_get, 1},
  {"CALL_nimb200_samples_summary", (DL_FUNC)&CALL_nimb200_samples_summary, 4},
  {"CALL_nimb200_history_enable", (DL_FUNC)&CALL_nimb200_history_enable, 2},
  {"CALL_nimb200_history", (DL_FUNC)&CALL_nimb200_history, 4},
  {"CALL_nimb200_clear", (DL_FUNC)&CALL_nimb200_clear, 1},
  {NULL, NULL, 0}};
void R_init_nimb200_R(DllInfo *dll) {
  R_registerRoutines(dll, NULL, call_methods, NULL, NULL);
  R_useDynamicSymbols(dll, FALSE);
}

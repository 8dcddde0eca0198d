/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model table; one translation unit per model (mbo_model_<name>.c). */
#include <string.h>
#include "mbo.h"
extern const mbo_model mbo_model_pendulum, mbo_model_spring_damper, mbo_model_robot6dof, mbo_model_fourbar, mbo_model_rope6, mbo_model_three_springs, mbo_model_fourbar_planar, mbo_model_free_couple, mbo_model_furuta, mbo_model_planar_pendulum, mbo_model_rope3e,
    mbo_model_rope6c, mbo_model_rope3ec, mbo_model_rope20, mbo_model_rope50, mbo_model_rope20e, mbo_model_rope50e;
static const mbo_model* const MODELS[] = {&mbo_model_pendulum, &mbo_model_spring_damper, &mbo_model_robot6dof,
                                          &mbo_model_fourbar, &mbo_model_rope6, &mbo_model_three_springs,
                                          &mbo_model_fourbar_planar, &mbo_model_free_couple, &mbo_model_furuta, &mbo_model_planar_pendulum, &mbo_model_rope3e,
                                          &mbo_model_rope6c, &mbo_model_rope3ec, &mbo_model_rope20, &mbo_model_rope50, &mbo_model_rope20e, &mbo_model_rope50e};

const mbo_model* mbo_find_model(const char* name) {
  for (unsigned i = 0; i < sizeof(MODELS) / sizeof(MODELS[0]); i++)
    if (strcmp(MODELS[i]->name, name) == 0) return MODELS[i];
  return 0;
}
int mbo_model_dims(const char* name, int* n, int* m, int* nnz) {
  const mbo_model* md = mbo_find_model(name);
  if (!md) return -1;
  *n = md->n; *m = md->m; *nnz = md->nnz;
  return 0;
}
int mbo_eval(const char* name, int what, const double* u, const double* p, double t, const double* tb, double* out) {
  const mbo_model* md = mbo_find_model(name);
  if (!md) return -1;
  mbo_fn f = (what == 0 ? md->rhs : what == 1 ? md->jac : what == 2 ? md->dT : md->observed);
  if (!f) return -2;        /* e.g. the loop-based rope models have no analytic Jacobian */
  f(u, p, t, tb, out);
  return 0;
}
int mbo_nobs(const char* name) {
  const mbo_model* md = mbo_find_model(name);
  return md ? md->nobs : -1;
}
int mbo_pattern(const char* name, int* jrow, int* jcol, int* mass) {
  const mbo_model* md = mbo_find_model(name);
  if (!md) return -1;
  memcpy(jrow, md->jrow, sizeof(int) * md->nnz);
  memcpy(jcol, md->jcol, sizeof(int) * md->nnz);
  memcpy(mass, md->mass, sizeof(int) * md->n);
  return 0;
}

/* Checks the product's generated structured factor/solve pair on the CPU (tests only; the
 * oracle's own integrator uses the dense pivoted LU above it, never these). */
#include <stdlib.h>
int mbo_factor_solve(const char* name, const double* u, const double* p, double t, const double* tb, double c, double* b) {
  const mbo_model* md = mbo_find_model(name);
  if (!md || !md->factor) return -1;
  double* LU = (double*)calloc(md->nlu > 0 ? md->nlu : 1, sizeof(double));
  md->factor(u, p, t, tb, c, LU);
  md->solve(LU, c, p, b);
  free(LU);
  return 0;
}

/* largest |1/pivot| the product's generated factor() stored for this state (runtime pivot guard of the kernels) */
double mbo_pivmax(const char* name, const double* u, const double* p, double t, const double* tb, double c) {
  const mbo_model* md = mbo_find_model(name);
  if (!md || !md->factor || !md->pivmax) return -1.0;
  double* LU = (double*)calloc(md->nlu > 0 ? md->nlu : 1, sizeof(double));
  md->factor(u, p, t, tb, c, LU);
  const double g = md->pivmax(LU);
  free(LU);
  return g;
}

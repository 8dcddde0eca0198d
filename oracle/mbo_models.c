/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model table: the generated right-hand sides (csrc/generated/mbm_*.h, emitted from the
 * restated reference equations) compiled as plain C. */
#include <string.h>
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_pendulum.h"
#include "../multibody.jl_b200/csrc/generated/mbm_spring_damper.h"
#include "../multibody.jl_b200/csrc/generated/mbm_robot6dof.h"
#ifdef MBO_HAVE_FOURBAR
#include "../multibody.jl_b200/csrc/generated/mbm_fourbar.h"
#endif
#ifdef MBO_HAVE_ROPE
#include "../multibody.jl_b200/csrc/generated/mbm_rope6.h"
#endif

#define ENTRY(X) {#X, MBM_##X##_N, MBM_##X##_M, MBM_##X##_NNZ, mbm_##X##_jrow, mbm_##X##_jcol, mbm_##X##_mass, \
                  mbm_##X##_rhs, mbm_##X##_jac, mbm_##X##_dT, MBM_##X##_NLU, mbm_##X##_factor, mbm_##X##_solve}
static const mbo_model MODELS[] = {
  ENTRY(pendulum), ENTRY(spring_damper), ENTRY(robot6dof),
#ifdef MBO_HAVE_FOURBAR
  ENTRY(fourbar),
#endif
#ifdef MBO_HAVE_ROPE
  ENTRY(rope6),
#endif
};

const mbo_model* mbo_find_model(const char* name) {
  for (unsigned i = 0; i < sizeof(MODELS) / sizeof(MODELS[0]); i++)
    if (strcmp(MODELS[i].name, name) == 0) return &MODELS[i];
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
  (what == 0 ? md->rhs : what == 1 ? md->jac : md->dT)(u, p, t, tb, out);
  return 0;
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
  if (!md) return -1;
  double* LU = (double*)calloc(md->nlu > 0 ? md->nlu : 1, sizeof(double));
  md->factor(u, p, t, tb, c, LU);
  md->solve(LU, c, p, b);
  free(LU);
  return 0;
}

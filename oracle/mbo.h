/* ORACLE -- test infrastructure only. Never linked into the product. */
#ifndef MBO_H
#define MBO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
enum { MBO_ALG_TSIT5 = 1, MBO_ALG_RODAS5P = 2, MBO_ALG_RODAS4 = 3, MBO_ALG_FBDF = 4 };
typedef struct {
  int32_t alg;        /* MBO_ALG_* */
  int32_t jac_mode;   /* 0 analytic (generated), 1 forward differences (`autodiff=false`, test/test_robot.jl:288) */
  double abstol, reltol;
  double dt0;         /* 0 = automatic */
  double dtmin, dtmax;/* 0 = default */
  int64_t maxiters;   /* 0 = 100000 */
  /* step-size dead band of OrdinaryDiffEq's PI controller (`step_accept_controller!`): when
   * qsteady_min <= q <= qsteady_max the step is kept (q = 1).  Explicit methods: (1, 1);
   * adaptive implicit / Rosenbrock methods: (1, 6/5).  0, 0 = the method's default. */
  double qsteady_min, qsteady_max;
  int32_t pow_mode;   /* 0 = OrdinaryDiffEq's `fastpower` (Float32 exp2 * approximate log2) in the PI controller, 1 = exact pow */
  int32_t reserved;
} mbo_opts;

typedef void (*mbo_fn)(const double* u, const double* p, double t, const double* tb, double* out);
typedef struct {
  const char* name;
  int n, m, nnz;
  const int *jrow, *jcol, *mass;
  mbo_fn rhs, jac, dT;
  int nlu;            /* slots of the product's generated structured LU (checked in tests only) */
  void (*factor)(const double* u, const double* p, double t, const double* tb, double c, double* LU);
  void (*solve)(const double* LU, double c, const double* p, double* b);
  int nobs;
  mbo_fn observed;
  double (*pivmax)(const double* LU);   /* largest stored |1/pivot| of the product's statically ordered LU (tests only) */
} mbo_model;

const mbo_model* mbo_find_model(const char* name);
int mbo_model_dims(const char* name, int* n, int* m, int* nnz);
int mbo_eval(const char* name, int what /*0 rhs,1 jac(nnz),2 dT,3 observed*/, const double* u, const double* p, double t, const double* tb, double* out);
int mbo_factor_solve(const char* name, const double* u, const double* p, double t, const double* tb, double c, double* b);
int mbo_pattern(const char* name, int* jrow, int* jcol, int* mass);
double mbo_pivmax(const char* name, const double* u, const double* p, double t, const double* tb, double c);
int mbo_solve_one(const mbo_model* md, const mbo_opts* o, const double* u0, const double* p, const double* tb,
                  double t0, double t1, const double* saveat, int nsave, double* out_u, double* out_tfinal, int32_t* stats);
int mbo_solve_ensemble(const char* model, const mbo_opts* o, int64_t N, const double* u0, const double* p, const double* tb,
                       double t0, double t1, const double* saveat, int nsave, double* out_u, double* out_tfinal,
                       int32_t* retcode, int32_t* stats, int nthreads);
#ifdef __cplusplus
}
#endif
#endif

/* ORACLE -- test infrastructure only. Never linked into the product (see oracle/README.md).
 *
 * CPU restatement of the reference's hot loop for one trajectory and its ensemble
 * fan-out: `solve(prob, Tsit5()/Rodas5P(); abstol, reltol, saveat)` as called at
 * /root/reference/test/test_robot.jl:288, docs/src/examples/pendulum.md:68 and
 * SURVEY.md section 3.2/3.3.  The arithmetic lives in OrdinaryDiffEq [un-vendored;
 * Project.toml:34-55]; its algorithm is restated from SURVEY.md Appendix C:
 *   - C.1 Tsit5 (7 stages, FSAL, 4th-order free interpolant)
 *   - C.2 Rodas5P (8 stages, W = M/(gamma h) - J, dense LU with partial pivoting,
 *         3-vector dense output)
 *   - C.3 PI step controller (beta1 = 7/50, beta2 = 2/25, gamma = 0.9, qmin = 0.2,
 *         qmax = 10, qoldinit = 1e-4), RMS error norm scaled by
 *         abstol + max(|u0|,|u1|) reltol, Hairer initial step, saveat by dense output.
 * Parity: the Tsit5 path is pinned by reference-run outputs (test/test_worldforces.jl:136,180,
 * reproduced to 6.5e-12 / 4.7e-10, tests/test_oracle.py).  The Rosenbrock path is PARITY
 * UNPINNED below 3.5e-4 (Julia is absent here; the one Rodas5P reference-run output,
 * test/test_orientation_getters.jl:69,86, is met at the reference's own tolerance only); it is
 * held instead by scipy Radau/DOP853 at rtol 1e-12, analytic solutions and convergence order.
 */
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#include "mbo_tableaus.h"
#include "mbo.h"

/* SciMLBase.ReturnCode values */
enum { RC_SUCCESS = 1, RC_DTNAN = 3, RC_MAXITERS = 4, RC_DTLESSTHANMIN = 5, RC_UNSTABLE = 6, RC_INITIALFAILURE = 7 };

/* Rosenbrock tableau in the transformed (Hairer-Wanner) form, padded to 8 stages */
typedef struct {
  int nst, nexp;          /* stages; stages 0..nexp-1 use rows of A, the rest add the previous k */
  int order;
  double gamma;
  double A[8][7], C[8][7], CT[8], D[8], H[3][8];
} rosen_tab;
static void rosen_get(int alg, rosen_tab* T) {
  memset(T, 0, sizeof(*T));
  if (alg == MBO_ALG_RODAS4) {
    T->nst = 6; T->nexp = 5; T->order = 4; T->gamma = R4_GAMMA;
    for (int i = 0; i < 5; i++) for (int j = 0; j < 4; j++) T->A[i][j] = R4_A[i][j];
    for (int i = 0; i < 6; i++) { for (int j = 0; j < 5; j++) T->C[i][j] = R4_C[i][j]; T->CT[i] = R4_CT[i]; T->D[i] = R4_D[i]; }
    for (int r = 0; r < 2; r++) for (int j = 0; j < 6; j++) T->H[r][j] = R4_H[r][j];
  } else {
    T->nst = 8; T->nexp = 6; T->order = 5; T->gamma = RD_GAMMA;
    for (int i = 0; i < 6; i++) for (int j = 0; j < 5; j++) T->A[i][j] = RD_A[i][j];
    for (int i = 0; i < 8; i++) { for (int j = 0; j < 7; j++) T->C[i][j] = RD_C[i][j]; T->CT[i] = RD_CT[i]; T->D[i] = RD_D[i]; }
    for (int r = 0; r < 3; r++) for (int j = 0; j < 8; j++) T->H[r][j] = RD_H[r][j];
  }
}

typedef struct {
  const mbo_model* md;
  const double* p;
  const double* tb;
  int n;
  int64_t nf, njac, nsolve, naccept, nreject;
} ctx_t;

static void f_eval(ctx_t* c, const double* u, double t, double* du) {
  c->md->rhs(u, c->p, t, c->tb, du);
  c->nf++;
}

/* OrdinaryDiffEq's PI controller raises EEst and qold to beta1 / beta2 with `fastpower`
 * (FastPower.jl; DiffEqBase.fastpow before it): Float32 exp2 of y times an approximate log2
 * (rational fit on the significand, D. Goldberg, "Fast approximate logarithms III", table 2 line 8),
 * so the step-size factors carry a relative error of ~1e-5 -- reproduced here because step sequences,
 * not just converged solutions, have to match the reference's runs. */
static float fastlog2f(float x) {
  const float a = 0.338953f, b = 2.198599f, c = 1.523692f;
  uint32_t ux1i; memcpy(&ux1i, &x, 4);
  uint32_t ex = (ux1i & 0x7F800000u) >> 23;
  uint32_t greater = ux1i & 0x00400000u;
  uint32_t ux2i; float signif, fexp;
  if (greater) { ux2i = (ux1i & 0x007FFFFFu) | 0x3f000000u; memcpy(&signif, &ux2i, 4); fexp = (float)ex - 126.0f; }
  else         { ux2i = (ux1i & 0x007FFFFFu) | 0x3f800000u; memcpy(&signif, &ux2i, 4); fexp = (float)ex - 127.0f; }
  signif = signif - 1.0f;
  return fexp + signif * (a * signif + b) / (signif + c);
}
static double fastpower(double x, double y) {
  if (x == 0.0) return 0.0;
  if (isinf(x) && isinf(y)) return INFINITY;
  /* exp2 evaluated in double and rounded to Float32: the same bits on the CPU and on the GPU
   * (libm's and CUDA's exp2f differ by an ulp, which would flip dead-band decisions) */
  return (double)(float)exp2((double)((float)y * fastlog2f((float)x)));
}
static double ctl_pow(const mbo_opts* o, double x, double y) { return o->pow_mode == 1 ? pow(x, y) : fastpower(x, y); }

static double rms_scaled(int n, const double* e, const double* ua, const double* ub, double atol, double rtol) {
  double s = 0.0;
  for (int i = 0; i < n; i++) {
    double sc = atol + fmax(fabs(ua[i]), fabs(ub[i])) * rtol;
    double r = e[i] / sc;
    s += r * r;
  }
  return sqrt(s / n);
}

/* OrdinaryDiffEq `ode_determine_initdt` (Hairer-Norsett-Wanner II.4), order-5 methods.
 * Algebraic rows (mass 0) carry no derivative information and are left out of d1/d2. */
static double initial_dt(ctx_t* c, const double* u0, double t0, double t1, const mbo_opts* o, int order, double* w1, double* w2, double* w3) {
  int n = c->n;
  const int* mass = c->md->mass;
  double dtmax = (o->dtmax > 0) ? o->dtmax : (t1 - t0);
  double* f0 = w1; double* u1 = w2; double* f1 = w3;
  f_eval(c, u0, t0, f0);
  double d0 = 0, d1 = 0;
  for (int i = 0; i < n; i++) {
    double sk = o->abstol + fabs(u0[i]) * o->reltol;
    d0 += (u0[i] / sk) * (u0[i] / sk);
    if (mass[i]) d1 += (f0[i] / sk) * (f0[i] / sk);
  }
  d0 = sqrt(d0 / n); d1 = sqrt(d1 / n);
  double dt0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
  dt0 = fmin(dt0, dtmax);
  for (int i = 0; i < n; i++) u1[i] = u0[i] + (mass[i] ? dt0 * f0[i] : 0.0);
  f_eval(c, u1, t0 + dt0, f1);
  double d2 = 0;
  for (int i = 0; i < n; i++) {
    double sk = o->abstol + fabs(u0[i]) * o->reltol;
    if (mass[i]) { double r = (f1[i] - f0[i]) / sk; d2 += r * r; }
  }
  d2 = sqrt(d2 / n) / dt0;
  double dm = fmax(d1, d2);
  double dt1 = (dm <= 1e-15) ? fmax(1e-6, dt0 * 1e-3) : pow(10.0, -(2.0 + log10(dm)) / (double)order);
  return fmin(fmin(100.0 * dt0, dt1), dtmax);
}

/* dense LU with partial pivoting, row-major A[n*n], in place */
static int lu_factor(int n, double* A, int* piv) {
  for (int k = 0; k < n; k++) {
    int pk = k; double mx = fabs(A[k * n + k]);
    for (int i = k + 1; i < n; i++) { double v = fabs(A[i * n + k]); if (v > mx) { mx = v; pk = i; } }
    piv[k] = pk;
    if (mx == 0.0 || !isfinite(mx)) return -1;
    if (pk != k) for (int j = 0; j < n; j++) { double tmp = A[k * n + j]; A[k * n + j] = A[pk * n + j]; A[pk * n + j] = tmp; }
    double inv = 1.0 / A[k * n + k];
    for (int i = k + 1; i < n; i++) {
      double l = A[i * n + k] * inv;
      A[i * n + k] = l;
      if (l != 0.0) for (int j = k + 1; j < n; j++) A[i * n + j] -= l * A[k * n + j];
    }
  }
  return 0;
}
static void lu_solve_dense(int n, const double* A, const int* piv, double* b) {
  /* rows of L were swapped as whole rows during factorisation (LAPACK convention), so
   * all interchanges are applied to b before the forward substitution */
  for (int k = 0; k < n; k++) {
    int pk = piv[k];
    if (pk != k) { double t = b[k]; b[k] = b[pk]; b[pk] = t; }
  }
  for (int k = 0; k < n; k++)
    for (int i = k + 1; i < n; i++) b[i] -= A[i * n + k] * b[k];
  for (int i = n - 1; i >= 0; i--) {
    double s = b[i];
    for (int j = i + 1; j < n; j++) s -= A[i * n + j] * b[j];
    b[i] = s / A[i * n + i];
  }
}

/* J dense row-major; analytic (generated) or forward differences as `autodiff=false` does */
static void jac_dense(ctx_t* c, const double* u, double t, const double* f0, int mode, double* J, double* wu, double* wf, double* Jnz) {
  int n = c->n;
  memset(J, 0, sizeof(double) * n * n);
  if (mode == 0) {
    c->md->jac(u, c->p, t, c->tb, Jnz);
    for (int k = 0; k < c->md->nnz; k++) J[c->md->jrow[k] * n + c->md->jcol[k]] = Jnz[k];
  } else {
    memcpy(wu, u, sizeof(double) * n);
    const double rel = sqrt(DBL_EPSILON);
    for (int j = 0; j < n; j++) {
      double eps = fmax(rel * fabs(u[j]), rel);
      wu[j] = u[j] + eps;
      f_eval(c, wu, t, wf);
      wu[j] = u[j];
      for (int i = 0; i < n; i++) J[i * n + j] = (wf[i] - f0[i]) / eps;
    }
  }
  c->njac++;
}

static void dT_eval(ctx_t* c, const double* u, double t, const double* f0, int mode, double* dT, double* wf) {
  int n = c->n;
  if (mode == 0) { c->md->dT(u, c->p, t, c->tb, dT); return; }
  double eps = sqrt(DBL_EPSILON) * fmax(fabs(t), 1.0);
  f_eval(c, u, t + eps, wf);
  for (int i = 0; i < n; i++) dT[i] = (wf[i] - f0[i]) / eps;
}

#include "mbo_bdf.inc"

int mbo_solve_one(const mbo_model* md, const mbo_opts* o, const double* u0, const double* p, const double* tb,
                  double t0, double t1, const double* saveat, int nsave, double* out_u /* [nsave][n] */,
                  double* out_tfinal, int32_t* stats /* [5] */) {
  if (o->alg == MBO_ALG_FBDF) return bdf_solve_one(md, o, u0, p, tb, t0, t1, saveat, nsave, out_u, out_tfinal, stats);
  const int n = md->n;
  ctx_t c = {md, p, tb, n, 0, 0, 0, 0, 0};
  const int rosen = (o->alg == MBO_ALG_RODAS5P || o->alg == MBO_ALG_RODAS4);
  rosen_tab T;
  rosen_get(o->alg, &T);
  const int order = rosen ? T.order : 5;
  const int nk = rosen ? T.nst : 7;
  size_t nd = (size_t)n * (nk + 12) + (size_t)n * n * 2 + md->nnz + 16;
  double* mem = (double*)calloc(nd, sizeof(double));
  int* piv = (int*)calloc(n, sizeof(int));
  double* k = mem;                  /* k[nk][n] */
  double* u = k + (size_t)nk * n;   /* current state */
  double* un = u + n;               /* candidate */
  double* us = un + n;              /* stage state */
  double* du = us + n;
  double* rhs = du + n;
  double* dT = rhs + n;
  double* w1 = dT + n; double* w2 = w1 + n; double* w3 = w2 + n;
  double* err = w3 + n;
  double* fsal = err + n;
  double* tmp = fsal + n;
  double* J = tmp + n;
  double* W = J + (size_t)n * n;
  double* Jnz = W + (size_t)n * n;
  const int* mass = md->mass;

  memcpy(u, u0, sizeof(double) * n);
  int is = 0;
  double t = t0;
  double dtmax = (o->dtmax > 0) ? o->dtmax : (t1 - t0);
  /* Consistent initialisation of the algebraic states of a mass-matrix DAE (what MTK's
   * initialisation problem / OrdinaryDiffEq's BrownFullBasicInit do before the first step,
   * test/test_fourbar.jl:44,70): differential states are kept, the algebraic equations
   * 0 = f_alg(u) are solved by Newton for the algebraic states.  One Newton step is a solve
   * with W = c M - J at c = 1e20, which pins the differential rows. */
  int init_fail = 0;
  {
    int is_dae = 0;
    for (int i = 0; i < n; i++) if (!mass[i]) is_dae = 1;
    if (is_dae && rosen) {
      const double tol_init = fmin(o->abstol, 1e-10);
      init_fail = 1;
      for (int it = 0; it <= 10; it++) {
        f_eval(&c, u, t0, du);
        double r = 0.0;
        for (int i = 0; i < n; i++) if (!mass[i]) r = fmax(r, fabs(du[i]));
        if (!(r > tol_init)) { init_fail = 0; break; }
        if (it == 10) break;
        jac_dense(&c, u, t0, du, 0, J, w1, w2, Jnz);
        c.njac--;
        for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) W[i * n + j] = -J[i * n + j] + ((i == j && mass[i]) ? 1e20 : 0.0);
        if (lu_factor(n, W, piv) != 0) break;
        for (int i = 0; i < n; i++) rhs[i] = mass[i] ? 0.0 : du[i];
        lu_solve_dense(n, W, piv, rhs);
        for (int i = 0; i < n; i++) if (!mass[i]) u[i] += rhs[i];
      }
    }
  }
  while (is < nsave && saveat[is] <= t0) { if (saveat[is] == t0) memcpy(out_u + (size_t)is * n, u, sizeof(double) * n); is++; }
  double dt = (o->dt0 > 0) ? o->dt0 : initial_dt(&c, u, t0, t1, o, order, w1, w2, w3);
  const double beta1 = 7.0 / (10.0 * order), beta2 = 2.0 / (5.0 * order), gam = 0.9, qmin = 0.2, qmax = 10.0;
  double qold = 1e-4;
  double qs_min = o->qsteady_min, qs_max = o->qsteady_max;
  if (qs_min == 0.0 && qs_max == 0.0) { qs_min = 1.0; qs_max = rosen ? 1.2 : 1.0; }
  int64_t maxiters = o->maxiters > 0 ? o->maxiters : 100000;
  int64_t iter = 0;
  int rc = RC_SUCCESS;
  int have_fsal = 0;
  const int trace = getenv("MBO_TRACE") != NULL;

  if (init_fail) rc = RC_INITIALFAILURE;          /* inconsistent algebraic states: do not integrate */
  while (t < t1 && !init_fail) {
    if (++iter > maxiters) { rc = RC_MAXITERS; break; }
    if (!isfinite(dt)) { rc = RC_DTNAN; break; }
    if (dt < fmax(o->dtmin, 16.0 * DBL_EPSILON * fmax(fabs(t), 1e-300))) { rc = RC_DTLESSTHANMIN; break; }
    int last = 0;
    if (t + dt >= t1 - 100.0 * DBL_EPSILON * fabs(t1)) { dt = t1 - t; last = 1; }
    double EEst;
    if (!rosen) {
      /* ---- Tsit5 (Appendix C.1) ---- */
      if (!have_fsal) { f_eval(&c, u, t, fsal); have_fsal = 1; }
      memcpy(k, fsal, sizeof(double) * n);
      for (int s = 1; s < 7; s++) {
        for (int i = 0; i < n; i++) {
          double a = 0.0;
          for (int j = 0; j < s; j++) a += TS_A[s][j] * k[(size_t)j * n + i];
          us[i] = u[i] + dt * a;
        }
        f_eval(&c, us, t + TS_C[s] * dt, k + (size_t)s * n);
      }
      memcpy(un, us, sizeof(double) * n);             /* stage 7 state is the new solution */
      for (int i = 0; i < n; i++) {
        double e = 0.0;
        for (int j = 0; j < 7; j++) e += TS_BT[j] * k[(size_t)j * n + i];
        err[i] = dt * e;
      }
    } else {
      /* ---- Rodas5P (Appendix C.2) ---- */
      f_eval(&c, u, t, du);
      /* models without an analytic Jacobian (loop-based ropes) are differenced; they are autonomous, df/dt = 0 analytically */
      jac_dense(&c, u, t, du, md->jac ? o->jac_mode : 1, J, w1, w2, Jnz);
      dT_eval(&c, u, t, du, md->jac ? o->jac_mode : 0, dT, w2);
      double cg = 1.0 / (dt * T.gamma);
      for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) W[i * n + j] = -J[i * n + j] + ((i == j && mass[i]) ? cg : 0.0);
      if (lu_factor(n, W, piv) != 0) { rc = RC_UNSTABLE; break; }
      memcpy(us, u, sizeof(double) * n);
      for (int s = 0; s < T.nst; s++) {
        if (s > 0) {
          if (s < T.nexp) {
            for (int i = 0; i < n; i++) {
              double a = 0.0;
              for (int j = 0; j < s; j++) a += T.A[s][j] * k[(size_t)j * n + i];
              us[i] = u[i] + a;
            }
          } else {
            for (int i = 0; i < n; i++) us[i] += k[(size_t)(s - 1) * n + i];
          }
          f_eval(&c, us, t + T.CT[s] * dt, du);
        }
        for (int i = 0; i < n; i++) {
          double a = 0.0;
          if (mass[i]) { for (int j = 0; j < s; j++) a += T.C[s][j] * k[(size_t)j * n + i]; a /= dt; }
          rhs[i] = du[i] + dt * T.D[s] * dT[i] + a;
        }
        lu_solve_dense(n, W, piv, rhs); c.nsolve++;
        memcpy(k + (size_t)s * n, rhs, sizeof(double) * n);
      }
      for (int i = 0; i < n; i++) { un[i] = us[i] + k[(size_t)(T.nst - 1) * n + i]; err[i] = k[(size_t)(T.nst - 1) * n + i]; }
    }
    EEst = rms_scaled(n, err, u, un, o->abstol, o->reltol);
    if (!isfinite(EEst)) { rc = RC_DTNAN; break; }
    /* PI controller (Appendix C.3) */
    double q, q11 = 0.0;
    if (EEst == 0.0) q = 1.0 / qmax;
    else {
      q11 = ctl_pow(o, EEst, beta1);
      q = q11 / ctl_pow(o, qold, beta2);
      q = fmax(1.0 / qmax, fmin(1.0 / qmin, q / gam));
    }
    if (trace) fprintf(stderr, "step t=%.17g dt=%.17g EEst=%.6e q=%.6f %s\n", t, dt, EEst, q, EEst <= 1.0 ? "acc" : "REJ");
    if (EEst <= 1.0) {
      c.naccept++;
      double tn = last ? t1 : t + dt;
      /* saveat through the dense output of this step */
      while (is < nsave && saveat[is] <= tn) {
        double ts = saveat[is];
        double* o_u = out_u + (size_t)is * n;
        if (ts == tn) memcpy(o_u, un, sizeof(double) * n);
        else {
          double th = (ts - t) / dt;
          if (!rosen) {
            double b[7];
            for (int j = 0; j < 7; j++) b[j] = th * (TS_R[j][0] + th * (TS_R[j][1] + th * (TS_R[j][2] + th * TS_R[j][3])));
            for (int i = 0; i < n; i++) {
              double a = 0.0;
              for (int j = 0; j < 7; j++) a += b[j] * k[(size_t)j * n + i];
              o_u[i] = u[i] + dt * a;
            }
          } else {
            for (int i = 0; i < n; i++) {
              double K1 = 0, K2 = 0, K3 = 0;
              for (int j = 0; j < T.nst; j++) { double kj = k[(size_t)j * n + i]; K1 += T.H[0][j] * kj; K2 += T.H[1][j] * kj; K3 += T.H[2][j] * kj; }
              o_u[i] = (1.0 - th) * u[i] + th * (un[i] + (1.0 - th) * (K1 + th * (K2 + th * K3)));
            }
          }
        }
        is++;
      }
      if (!rosen) memcpy(fsal, k + (size_t)6 * n, sizeof(double) * n);
      memcpy(u, un, sizeof(double) * n);
      t = tn;
      qold = fmax(EEst, 1e-4);
      if (qs_min <= q && q <= qs_max) q = 1.0;          /* dead band of step_accept_controller! */
      dt = fmin(dt / q, dtmax);
    } else {
      c.nreject++;
      dt = dt / fmin(1.0 / qmin, q11 / gam);
    }
  }
  *out_tfinal = t;
  stats[0] = (int32_t)c.naccept; stats[1] = (int32_t)c.nreject; stats[2] = (int32_t)c.nf;
  stats[3] = (int32_t)c.njac; stats[4] = (int32_t)c.nsolve;
  (void)tmp;
  free(mem); free(piv);
  return rc;
}

/* Ensemble fan-out (SciMLBase EnsembleThreads equivalent): SoA inputs/outputs in the
 * same layout as the product's C ABI (include/mbb200.h).  pthreads, dynamic chunks. */
#include <pthread.h>
#include <stdatomic.h>
typedef struct {
  const mbo_model* md; const mbo_opts* o; int64_t N; const double* u0; const double* p; const double* tb;
  double t0, t1; const double* saveat; int nsave; double* out_u; double* out_tfinal; int32_t* retcode; int32_t* stats;
  atomic_llong next;
} ens_t;

static void* ens_worker(void* arg) {
  ens_t* e = (ens_t*)arg;
  const int n = e->md->n, m = e->md->m, nsave = e->nsave;
  const int64_t N = e->N;
  double* ul = (double*)malloc(sizeof(double) * (n + (m > 0 ? m : 1)));
  double* pl = ul + n;
  double* ol = (double*)malloc(sizeof(double) * (size_t)(nsave > 0 ? nsave : 1) * n);
  for (;;) {
    int64_t b = atomic_fetch_add(&e->next, 4);
    if (b >= N) break;
    int64_t hi = b + 4 < N ? b + 4 : N;
    for (int64_t i = b; i < hi; i++) {
      for (int j = 0; j < n; j++) ul[j] = e->u0[(size_t)j * N + i];
      for (int j = 0; j < m; j++) pl[j] = e->p[(size_t)j * N + i];
      int32_t st[5]; double tf;
      for (size_t q = 0; q < (size_t)nsave * n; q++) ol[q] = NAN;
      int rc = mbo_solve_one(e->md, e->o, ul, pl, e->tb, e->t0, e->t1, e->saveat, nsave, ol, &tf, st);
      e->retcode[i] = rc; e->out_tfinal[i] = tf;
      for (int s = 0; s < 5; s++) e->stats[(size_t)s * N + i] = st[s];
      for (int s = 0; s < nsave; s++) for (int j = 0; j < n; j++) e->out_u[((size_t)s * n + j) * N + i] = ol[(size_t)s * n + j];
    }
  }
  free(ul); free(ol);
  return 0;
}

int mbo_solve_ensemble(const char* model, const mbo_opts* o, int64_t N, const double* u0 /* [n][N] */,
                       const double* p /* [m][N] */, const double* tb, double t0, double t1,
                       const double* saveat, int nsave, double* out_u /* [nsave][n][N] */,
                       double* out_tfinal, int32_t* retcode, int32_t* stats /* [5][N] */, int nthreads) {
  const mbo_model* md = mbo_find_model(model);
  if (!md) return -1;
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  ens_t e = {md, o, N, u0, p, tb, t0, t1, saveat, nsave, out_u, out_tfinal, retcode, stats, 0};
  pthread_t th[256];
  for (int k = 1; k < nthreads; k++) pthread_create(&th[k], 0, ens_worker, &e);
  ens_worker(&e);
  for (int k = 1; k < nthreads; k++) pthread_join(th[k], 0);
  return 0;
}

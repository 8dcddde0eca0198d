// Variable-order BDF (orders 1-5) ensemble kernel for `FBDF()`, the solver of the reference's loop examples
// (/root/reference/test/test_fourbar.jl:45,71; docs/src/examples/kinematic_loops.md:80,152,190).
//
// Same algorithm, operation for operation, as oracle/mbo_bdf.inc: the NDF formulation of Shampine & Reichelt (quasi-constant
// step, differences D_0..D_{k+2}, kappa-corrected coefficients) for mass-matrix problems M y' = f, simplified Newton with
// the Jacobian taken at the predictor.  NOT step-identical to OrdinaryDiffEqBDF's FBDF (un-vendored; see the oracle header
// and DESIGN.md section 4).  The Newton matrix M / c - J, c = h / alpha_k, has the form of the Rosenbrock W, so the
// generated fused `factor` (W-build + structured LU) and `solve` are reused as they are.
// One trajectory per thread, global work dispenser, block-lockstep stepping (see mb_ensemble.cuh).
#pragma once
#include "mb_ensemble.cuh"

namespace mbk {

static __constant__ double cBDF_KAPPA[7] = {0.0, -0.1850, -1.0 / 9.0, -0.0823, -0.0415, 0.0, 0.0};
constexpr int BDF_MAXORD = 5, BDF_NEWTON_MAXIT = 4;

// D_0..D_order <- (R(factor) U)' D_0..D_order  (rescaling of the differences when the step changes by `factor`)
template <int N>
__device__ __noinline__ void bdf_change_D(double (*D)[N], int order, double factor) {
  double R[6][6], U[6][6], RU[6][6];
  for (int j = 0; j <= order; j++) { R[0][j] = 1.0; U[0][j] = 1.0; }
  for (int i = 1; i <= order; i++)
    for (int j = 0; j <= order; j++) {
      const double mr = (j == 0) ? 0.0 : ((double)(i - 1) - factor * (double)j) / (double)i;
      const double mu = (j == 0) ? 0.0 : ((double)(i - 1) - (double)j) / (double)i;
      R[i][j] = R[i - 1][j] * mr;
      U[i][j] = U[i - 1][j] * mu;
    }
  for (int i = 0; i <= order; i++)
    for (int j = 0; j <= order; j++) {
      double s = 0.0;
      for (int k = 0; k <= order; k++) s += R[i][k] * U[k][j];
      RU[i][j] = s;
    }
  for (int q = 0; q < N; q++) {
    double col[6], out[6];
    for (int i = 0; i <= order; i++) col[i] = D[i][q];
    for (int j = 0; j <= order; j++) {
      double s = 0.0;
      for (int i = 0; i <= order; i++) s += RU[i][j] * col[i];
      out[j] = s;
    }
    for (int j = 0; j <= order; j++) D[j][q] = out[j];
  }
}

template <int N>
__device__ __forceinline__ double bdf_norm(const double* x, const double* scale) {
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < N; i++) { const double r = x[i] / scale[i]; s += r * r; }
  return sqrt(s / N);
}

// M supplies: N, NP, NLU, IS_ODE, rhs(), factor(), solve(), pivmax(), isdiff().
template <class M, int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) bdf_kernel(const MbArgs a) {
  constexpr int N = M::N;
  constexpr int NP = (M::NP > 0 ? M::NP : 1);
  constexpr int NLU = (M::NLU > 0 ? M::NLU : 1);
  constexpr double EPS = 2.220446049250313e-16;
  const double INF = __longlong_as_double(0x7ff0000000000000LL);

  double D[8][N], y[N], d[N], psi[N], r[N], scale[N], p[NP], LU[NLU];
  double gam[7], alpha[7], econst[7];
  gam[0] = 0.0;
  for (int k = 1; k <= 6; k++) gam[k] = gam[k - 1] + 1.0 / k;
  for (int k = 0; k <= 6; k++) {
    const double kap = cBDF_KAPPA[k];
    alpha[k] = (1.0 - kap) * gam[k];
    econst[k] = kap * gam[k] + 1.0 / (k + 1);
  }
  double t = 0.0, h = 0.0;
  long long traj = -1, iter = 0;
  int is = 0, order = 1, n_equal = 0, naccept = 0, nreject = 0, nf = 0, njac = 0, nsolve = 0;
  bool active = false, drained = false;
  const double dtmax = a.dtmax > 0.0 ? a.dtmax : (a.t1 - a.t0);
  const long long maxiters = a.maxiters > 0 ? a.maxiters : 100000;
  const double newton_tol = fmax(10.0 * EPS / a.reltol, fmin(0.03, sqrt(a.reltol)));
  const double* __restrict__ tb = a.tb;

  for (;;) {
    int rc = RC_DEFAULT;
    if (!active && !drained) {
      const unsigned long long idx = atomicAdd(a.queue, 1ull);
      if (idx >= (unsigned long long)a.N) {
        drained = true;
      } else {
        traj = (long long)idx;
#pragma unroll
        for (int i = 0; i < N; i++) y[i] = a.u0[(size_t)i * a.N + traj];
#pragma unroll
        for (int i = 0; i < M::NP; i++) p[i] = a.p[(size_t)i * a.N + traj];
        t = a.t0; is = 0; iter = 0; order = 1; n_equal = 0;
        naccept = nreject = nf = njac = nsolve = 0;
        bool init_fail = false;
        if (!M::IS_ODE) {
          const int k0 = dae_init<M>(y, p, tb, a.t0, a.abstol, LU);
          init_fail = k0 < 0;
          nf += k0 < 0 ? -k0 : k0;
        }
        while (is < a.n_save && a.saveat[is] <= t) {
          if (a.saveat[is] == t) {
#pragma unroll
            for (int i = 0; i < N; i++) a.out_u[((size_t)is * N + i) * a.N + traj] = y[i];
          }
          is++;
        }
        if (a.dt0 > 0.0) h = a.dt0;
        else { h = initial_dt<M>(y, p, tb, a.t0, dtmax, a.abstol, a.reltol, 1); nf += 2; }
        M::rhs(y, p, t, tb, r); nf++;
#pragma unroll
        for (int i = 0; i < N; i++) {
          D[0][i] = y[i];
          D[1][i] = M::isdiff(i) ? r[i] * h : 0.0;
#pragma unroll
          for (int k = 2; k < 8; k++) D[k][i] = 0.0;
        }
        active = true;
        if (!(t < a.t1)) rc = RC_SUCCESS;
        if (init_fail) rc = RC_INITIALFAILURE;
      }
    }
#ifdef MB_NO_LOCKSTEP
    if (!__any_sync(0xffffffffu, active)) break;
#else
    if (!__syncthreads_or(active)) break;
#endif
    if (active && rc == RC_DEFAULT) {
      ++iter;
      if (iter > maxiters) rc = RC_MAXITERS;
      else if (!isfinite(h)) rc = RC_DTNAN;
      else if (h < fmax(a.dtmin, 16.0 * EPS * fmax(fabs(t), 1e-300))) rc = RC_DTLESSTHANMIN;
    }
    if (active && rc == RC_DEFAULT) {
      // ---- one step attempt -----------------------------------------------------------------------------------
      if (h > dtmax) { bdf_change_D<N>(D, order, dtmax / h); h = dtmax; n_equal = 0; }
      double t_new = t + h;
      if (t_new >= a.t1 - 100.0 * EPS * fabs(a.t1)) {
        if (a.t1 - t != h) { bdf_change_D<N>(D, order, (a.t1 - t) / h); n_equal = 0; }
        h = a.t1 - t; t_new = a.t1;
      }
      for (int i = 0; i < N; i++) {                      // predictor
        double s = 0.0, ps = 0.0;
        for (int k = 0; k <= order; k++) s += D[k][i];
        for (int k = 1; k <= order; k++) ps += gam[k] * D[k][i];
        y[i] = s; psi[i] = ps / alpha[order];
        scale[i] = a.abstol + a.reltol * fabs(s);
        d[i] = 0.0;
      }
      const double cc = h / alpha[order];
      const double cW = 1.0 / cc;
      M::factor(y, p, t_new, tb, cW, LU);                // Newton matrix M / c - J at the predictor
      njac++;
      const bool badpiv = !(M::pivmax(LU) <= MB_PIVOT_GUARD);
      bool converged = false;
      int nit = 0;
      double dy_old = -1.0, rate = 0.0;
      if (!badpiv) {
        for (int k = 0; k < BDF_NEWTON_MAXIT; k++) {     // simplified Newton
          M::rhs(y, p, t_new, tb, r); nf++;
          bool finite = true;
#pragma unroll
          for (int i = 0; i < N; i++) {
            if (!isfinite(r[i])) finite = false;
            r[i] = r[i] - (M::isdiff(i) ? (psi[i] + d[i]) / cc : 0.0);
          }
          if (!finite) break;
          M::solve(LU, cW, p, r); nsolve++;
          const double dyn = bdf_norm<N>(r, scale);
          nit = k + 1;
          const bool tiny = dyn <= 1e-3 * newton_tol;    // see oracle/mbo_bdf.inc
          if (dy_old >= 0.0 && !tiny) {
            rate = dyn / dy_old;
            if (rate >= 1.0 || pow(rate, (double)(BDF_NEWTON_MAXIT - k)) / (1.0 - rate) * dyn > newton_tol) break;
          }
#pragma unroll
          for (int i = 0; i < N; i++) { y[i] += r[i]; d[i] += r[i]; }
          if (tiny || (dy_old >= 0.0 && rate / (1.0 - rate) * dyn < newton_tol)) { converged = true; break; }
          dy_old = dyn;
        }
      }
      if (badpiv) {
        rc = RC_UNSTABLE;
      } else if (!converged) {                           // the Jacobian is current: halve the step
        nreject++;
        bdf_change_D<N>(D, order, 0.5); h *= 0.5; n_equal = 0;
      } else {
        const double safety = 0.9 * (2.0 * BDF_NEWTON_MAXIT + 1.0) / (2.0 * BDF_NEWTON_MAXIT + nit);
#pragma unroll
        for (int i = 0; i < N; i++) { scale[i] = a.abstol + a.reltol * fabs(y[i]); r[i] = econst[order] * d[i]; }
        const double en = bdf_norm<N>(r, scale);
        if (!isfinite(en)) {
          rc = RC_DTNAN;
        } else if (en > 1.0) {
          nreject++;
          const double fac = fmax(0.2, safety * pow(en, -1.0 / (order + 1)));
          bdf_change_D<N>(D, order, fac); h *= fac; n_equal = 0;
        } else {
          naccept++; n_equal++;
          for (int i = 0; i < N; i++) {                  // update the differences
            D[order + 2][i] = d[i] - D[order + 1][i];
            D[order + 1][i] = d[i];
          }
          for (int k = order; k >= 0; k--)
            for (int i = 0; i < N; i++) D[k][i] += D[k + 1][i];
          while (is < a.n_save && a.saveat[is] <= t_new) {   // dense output: interpolating polynomial of the differences
            const double ts = a.saveat[is];
            double* __restrict__ o = a.out_u + (size_t)is * N * a.N + traj;
            if (ts == t_new) {
#pragma unroll
              for (int i = 0; i < N; i++) o[(size_t)i * a.N] = y[i];
            } else {
              double pk[BDF_MAXORD + 1], x = 1.0;
              for (int j = 1; j <= order; j++) { x *= (ts - (t_new - h * (j - 1))) / (h * j); pk[j] = x; }
              for (int i = 0; i < N; i++) {
                double s = D[0][i];
                for (int j = 1; j <= order; j++) s += D[j][i] * pk[j];
                o[(size_t)i * a.N] = s;
              }
            }
            is++;
          }
          t = t_new;
          if (!(t < a.t1)) rc = RC_SUCCESS;
          else if (n_equal >= order + 1) {               // order and step selection
            double em = INF, ep = INF;
            if (order > 1) { for (int i = 0; i < N; i++) r[i] = econst[order - 1] * D[order][i]; em = bdf_norm<N>(r, scale); }
            if (order < BDF_MAXORD) { for (int i = 0; i < N; i++) r[i] = econst[order + 1] * D[order + 2][i]; ep = bdf_norm<N>(r, scale); }
            const double fm = (em > 0.0 && isfinite(em)) ? pow(em, -1.0 / order) : (em == 0.0 ? INF : 0.0);
            const double f0 = en > 0.0 ? pow(en, -1.0 / (order + 1)) : INF;
            const double fp = (ep > 0.0 && isfinite(ep)) ? pow(ep, -1.0 / (order + 2)) : (ep == 0.0 ? INF : 0.0);
            int dord = 0; double fb = f0;
            if (fm > fb) { fb = fm; dord = -1; }
            if (fp > fb) { fb = fp; dord = 1; }
            order += dord;
            const double fac = fmin(10.0, safety * fb);
            bdf_change_D<N>(D, order, fac); h *= fac; n_equal = 0;
          }
        }
      }
    }
    if (active && rc != RC_DEFAULT) {
      a.out_tfinal[traj] = t;
      a.retcode[traj] = rc;
      a.stats[0 * a.N + traj] = naccept; a.stats[1 * a.N + traj] = nreject; a.stats[2 * a.N + traj] = nf;
      a.stats[3 * a.N + traj] = njac; a.stats[4 * a.N + traj] = nsolve;
      const double qnan = __longlong_as_double(0x7ff8000000000000LL);
      for (; is < a.n_save; is++)
        for (int i = 0; i < N; i++) a.out_u[((size_t)is * N + i) * a.N + traj] = qnan;
      mb_retire(a, traj);
      active = false;
    }
  }
}

}  // namespace mbk

// Ensemble integrator kernels: one trajectory per thread, warp-level work refill.
//
// Replaces, for ensembles, the reference's per-trajectory hot loop
//   solve(prob, Tsit5() | Rodas5P(); abstol, reltol, saveat)
// (/root/reference/test/test_robot.jl:288, docs/src/examples/pendulum.md:68)
// and the SciMLBase EnsembleThreads fan-out.
//
// Design (B200): FP64-pipe bound, no tensor cores (the per-trajectory systems are tiny
// and irregular).  Every lane owns one trajectory and runs the *same* straight-line step
// code, so lanes never diverge inside a step; lanes differ only in (t, dt).  A lane that
// finishes pulls the next trajectory index from a global counter, so a warp stays full
// until the ensemble is drained (no tail of idle lanes waiting for the slowest
// trajectory).  The stage loop is rolled (coefficients from __constant__ memory, uniform
// across the warp) so the generated rhs / factor / solve code is instantiated once and
// inlined: stage vectors with compile-time indices live in registers, only the k-history
// is in (L1-resident, lane-interleaved) local memory.
#pragma once
#include "mb_common.cuh"

namespace mbk {

// augmented Rodas5P a-table: rows 6,7 fold "u7 = u6 + k6", "u8 = u7 + k7" into sums
static __constant__ double cRD_A[8][7] = {
    {0, 0, 0, 0, 0, 0, 0},
    {3.0, 0, 0, 0, 0, 0, 0},
    {2.849394379747939, 0.45842242204463923, 0, 0, 0, 0, 0},
    {-6.954028509809101, 2.489845061869568, -10.358996098473584, 0, 0, 0, 0},
    {2.8029986275628964, 0.5072464736228206, -0.3988312541770524, -0.04721187230404641, 0, 0, 0},
    {-7.502846399306121, 2.561846144803919, -11.627539656261098, -0.18268767659942256, 0.030198172008377946, 0, 0},
    {-7.502846399306121, 2.561846144803919, -11.627539656261098, -0.18268767659942256, 0.030198172008377946, 1.0, 0},
    {-7.502846399306121, 2.561846144803919, -11.627539656261098, -0.18268767659942256, 0.030198172008377946, 1.0, 1.0}};
static __constant__ double cRD_C[8][7] = {
    {0, 0, 0, 0, 0, 0, 0},
    {-14.155112264123755, 0, 0, 0, 0, 0, 0},
    {-17.97296035885952, -2.859693295451294, 0, 0, 0, 0, 0},
    {147.12150275711716, -1.41221402718213, 71.68940251302358, 0, 0, 0, 0},
    {165.43517024871676, -0.4592823456491126, 42.90938336958603, -5.961986721573306, 0, 0, 0},
    {24.854864614690072, -3.0009227002832186, 47.4931110020768, 5.5814197821558125, -0.6610691825249471, 0, 0},
    {30.91273214028599, -3.1208243349937974, 77.79954646070892, 34.28646028294783, -19.097331116725623, -28.087943162872662, 0},
    {37.80277123390563, -3.2571969029072276, 112.26918849496327, 66.9347231244047, -40.06618937091002, -54.66780262877968, -9.48861652309627}};
static __constant__ double cRD_CT[8] = {0.0, 0.6358126895828704, 0.4095798393397535, 0.9769306725060716,
                                        0.4288403609558664, 1.0, 1.0, 1.0};
static __constant__ double cRD_D[8] = {0.21193756319429014, -0.42387512638858027, -0.3384627126235924,
                                       1.8046452872882734, 2.325825639765069, 0.0, 0.0, 0.0};
static __constant__ double cTS_A[7][6] = {
    {0, 0, 0, 0, 0, 0},
    {0.161, 0, 0, 0, 0, 0},
    {-0.008480655492356989, 0.335480655492357, 0, 0, 0, 0},
    {2.8971530571054935, -6.359448489975075, 4.3622954328695815, 0, 0, 0},
    {5.325864828439257, -11.748883564062828, 7.4955393428898365, -0.09249506636175525, 0, 0},
    {5.86145544294642, -12.92096931784711, 8.159367898576159, -0.071584973281401, -0.028269050394068383, 0},
    {0.09646076681806523, 0.01, 0.4798896504144996, 1.379008574103742, -3.290069515436081, 2.324710524099774}};
static __constant__ double cTS_C[7] = {0.0, 0.161, 0.327, 0.9, 0.9800255409045097, 1.0, 1.0};

// Rodas4 (Hairer & Wanner RODAS; OrdinaryDiffEq `Rodas4`, the solver most of the reference's
// tests use): 6 stages, augmented a-row 5 folds "u6 = u5 + k5"
static __constant__ double cR4_A[6][5] = {
    {0, 0, 0, 0, 0},
    {1.544, 0, 0, 0, 0},
    {0.9466785280815826, 0.2557011698983284, 0, 0, 0},
    {3.314825187068521, 2.896124015972201, 0.9986419139977817, 0, 0},
    {1.221224509226641, 6.019134481288629, 12.53708332932087, -0.6878860361058950, 0},
    {1.221224509226641, 6.019134481288629, 12.53708332932087, -0.6878860361058950, 1.0}};
static __constant__ double cR4_C[6][5] = {
    {0, 0, 0, 0, 0},
    {-5.6688, 0, 0, 0, 0},
    {-2.430093356833875, -0.2063599157091915, 0, 0, 0},
    {-0.1073529058151375, -9.594562251023355, -20.47028614809616, 0, 0},
    {7.496443313967647, -10.24680431464352, -33.99990352819905, 11.70890893206160, 0},
    {8.083246795921522, -7.981132988064893, -31.52159432874371, 16.31930543123136, -6.058818238834054}};
static __constant__ double cR4_CT[6] = {0.0, 0.386, 0.21, 0.63, 1.0, 1.0};
static __constant__ double cR4_D[6] = {0.25, -0.1043, 0.1035, -0.03620000000000023, 0.0, 0.0};
static __constant__ double cR4_H[3][6] = {
    {10.12623508344586, -7.487995877610167, -34.80091861555747, -7.992771707568823, 1.025137723295662, 0.0},
    {-0.6762803392801253, 6.087714651680015, 16.43084320892478, 24.76722511418386, -6.594389125716872, 0.0},
    {0, 0, 0, 0, 0, 0}};

// Rosenbrock tableaus as traits (all reads are warp-uniform constant-bank loads)
struct TabRodas5P {
  static constexpr int NST = 8, ORDER = 5;
  static constexpr double GAMMA = RD_GAMMA;
  static __device__ __forceinline__ double A(int s, int j) { return cRD_A[s][j]; }
  static __device__ __forceinline__ double C(int s, int j) { return cRD_C[s][j]; }
  static __device__ __forceinline__ double CT(int s) { return cRD_CT[s]; }
  static __device__ __forceinline__ double D(int s) { return cRD_D[s]; }
  static __device__ __forceinline__ double H(int r, int j) { return RD_H[r][j]; }
};
struct TabRodas4 {
  static constexpr int NST = 6, ORDER = 4;
  static constexpr double GAMMA = 0.25;
  static __device__ __forceinline__ double A(int s, int j) { return cR4_A[s][j]; }
  static __device__ __forceinline__ double C(int s, int j) { return cR4_C[s][j]; }
  static __device__ __forceinline__ double CT(int s) { return cR4_CT[s]; }
  static __device__ __forceinline__ double D(int s) { return cR4_D[s]; }
  static __device__ __forceinline__ double H(int r, int j) { return cR4_H[r][j]; }
};
template <int ALG> struct TabOf { using type = TabRodas5P; };
template <> struct TabOf<3> { using type = TabRodas4; };

template <int N>
__device__ __forceinline__ double scaled_rms(const double (&e)[N], const double (&ua)[N],
                                             const double (&ub)[N], double atol, double rtol) {
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < N; i++) {
    const double sc = atol + fmax(fabs(ua[i]), fabs(ub[i])) * rtol;
    const double r = e[i] / sc;
    s += r * r;
  }
  return sqrt(s / N);
}

// OrdinaryDiffEq `ode_determine_initdt` restated; same arithmetic
// as the CPU oracle so that both start from the same first step.
template <class M>
__device__ __noinline__ double initial_dt(const double (&u)[M::N], const double* p, const double* tb,
                                          double t0, double dtmax, double atol, double rtol, int order = 5) {
  constexpr int N = M::N;
  double f0[N], u1[N], f1[N];
  M::rhs(u, p, t0, tb, f0);
  double d0 = 0, d1 = 0;
#pragma unroll
  for (int i = 0; i < N; i++) {
    const double sk = atol + fabs(u[i]) * rtol;
    d0 += (u[i] / sk) * (u[i] / sk);
    if (M::isdiff(i)) d1 += (f0[i] / sk) * (f0[i] / sk);
  }
  d0 = sqrt(d0 / N);
  d1 = sqrt(d1 / N);
  double dt0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
  dt0 = fmin(dt0, dtmax);
#pragma unroll
  for (int i = 0; i < N; i++) u1[i] = u[i] + (M::isdiff(i) ? dt0 * f0[i] : 0.0);
  M::rhs(u1, p, t0 + dt0, tb, f1);
  double d2 = 0;
#pragma unroll
  for (int i = 0; i < N; i++) {
    const double sk = atol + fabs(u[i]) * rtol;
    if (M::isdiff(i)) {
      const double r = (f1[i] - f0[i]) / sk;
      d2 += r * r;
    }
  }
  d2 = sqrt(d2 / N) / dt0;
  const double dm = fmax(d1, d2);
  const double dt1 = (dm <= 1e-15) ? fmax(1e-6, dt0 * 1e-3) : pow(10.0, -(2.0 + log10(dm)) / (double)order);
  return fmin(fmin(100.0 * dt0, dt1), dtmax);
}

// Consistent initialisation of the algebraic states of a mass-matrix DAE, per trajectory
// (the step before the path: MTK's initialisation problem / BrownFullBasicInit in the
// reference's stack, test/test_fourbar.jl:44,70).  Differential states are kept; 0 = f_alg(u)
// is solved by Newton for the algebraic states.  A Newton step reuses the generated
// factor/solve pair with c = 1e20, which pins the differential rows of W = c M - J.
template <class M>
__device__ __noinline__ int dae_init(double* u, const double* p, const double* tb, double t0, double abstol,
                                     double* LU) {
  // returns the number of right-hand-side evaluations; NEGATIVE when the algebraic residual is
  // still above tolerance after 10 Newton steps (the caller retires the trajectory with
  // ReturnCode.InitialFailure instead of integrating from an inconsistent state)
  constexpr int N = M::N;
  const double tol = fmin(abstol, 1e-10);
  int nf = 0;
  bool ok = false;
  for (int it = 0; it <= 10; it++) {
    double f[N];
    M::rhs(u, p, t0, tb, f);
    nf++;
    double r = 0.0;
#pragma unroll
    for (int i = 0; i < N; i++) if (!M::isdiff(i)) r = fmax(r, fabs(f[i]));
    if (!(r > tol)) { ok = true; break; }
    if (it == 10) break;
    M::factor(u, p, t0, tb, 1e20, LU);
#pragma unroll
    for (int i = 0; i < N; i++) if (M::isdiff(i)) f[i] = 0.0;
    M::solve(LU, 1e20, p, f);
#pragma unroll
    for (int i = 0; i < N; i++) if (!M::isdiff(i)) u[i] += f[i];
  }
  return ok ? nf : -nf;
}

}  // namespace mbk

namespace mbk {

// M supplies: N, NP, NLU, HAS_DT, rhs(), dT(), factor(), solve(), isdiff().
// ALG: 1 = Tsit5, 2 = Rodas5P, 3 = Rodas4.  UNR: 1 -> stage loop fully unrolled (small models).
template <class M, int ALG, int BLOCK, int MINB, int UNR>
__global__ void __launch_bounds__(BLOCK, MINB) ensemble_kernel(const MbArgs a) {
  constexpr int N = M::N;
  constexpr int NP = (M::NP > 0 ? M::NP : 1);
  constexpr bool ROSEN = (ALG >= 2);
  using T = typename TabOf<ALG>::type;
  constexpr int NST = T::NST;
  constexpr int NK = ROSEN ? NST : 7;
  constexpr int ORDER = ROSEN ? T::ORDER : 5;
  constexpr double BETA1 = 7.0 / (10.0 * ORDER), BETA2 = 2.0 / (5.0 * ORDER);
  constexpr int NLU = ROSEN ? (M::NLU > 0 ? M::NLU : 1) : 1;
  constexpr int NDT = (ROSEN && M::HAS_DT) ? N : 1;
  constexpr unsigned FULL = 0xffffffffu;
  constexpr double EPS = 2.220446049250313e-16;

  double u[N], un[N], us[N], k[NK][N], p[NP], LU[NLU], dT[NDT];
  double t = 0.0, dt = 0.0, qold = QOLDINIT;
  long long traj = -1, iter = 0;
  int is = 0, naccept = 0, nreject = 0, nf = 0, njac = 0, nsolve = 0;
  bool active = false, drained = false;
  const double dtmax = a.dtmax > 0.0 ? a.dtmax : (a.t1 - a.t0);
  const long long maxiters = a.maxiters > 0 ? a.maxiters : 100000;
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
        for (int i = 0; i < N; i++) u[i] = a.u0[(size_t)i * a.N + traj];
#pragma unroll
        for (int i = 0; i < M::NP; i++) p[i] = a.p[(size_t)i * a.N + traj];
        t = a.t0; is = 0; iter = 0; qold = QOLDINIT;
        naccept = nreject = nf = njac = nsolve = 0;
        bool init_fail = false;
        if (ROSEN && !M::IS_ODE) {
          const int k0 = dae_init<M>(u, p, tb, a.t0, a.abstol, LU);
          init_fail = k0 < 0;
          nf += k0 < 0 ? -k0 : k0;
        }
        while (is < a.n_save && a.saveat[is] <= t) {
          if (a.saveat[is] == t) {
#pragma unroll
            for (int i = 0; i < N; i++) a.out_u[((size_t)is * N + i) * a.N + traj] = u[i];
          }
          is++;
        }
        if (a.dt0 > 0.0) dt = a.dt0;
        else { dt = initial_dt<M>(u, p, tb, a.t0, dtmax, a.abstol, a.reltol, ORDER); nf += 2; }
        if (!ROSEN) { M::rhs(u, p, t, tb, k[0]); nf++; }   // k[0] = f(u,t), kept across rejections
        active = true;
        if (!(t < a.t1)) rc = RC_SUCCESS;
        if (init_fail) rc = RC_INITIALFAILURE;
      }
    }
    // block-wide barrier + vote: keeps the warps of a CTA in step, so the (large) straight-line
    // instruction stream is fetched once per CTA per step instead of once per warp
#ifdef MB_NO_LOCKSTEP
    if (!__any_sync(FULL, active)) break;
#else
    if (!__syncthreads_or(active)) break;
#endif
    if (active && rc == RC_DEFAULT) {
      ++iter;
      if (iter > maxiters) rc = RC_MAXITERS;
      else if (!isfinite(dt)) rc = RC_DTNAN;
      else if (dt < fmax(a.dtmin, 16.0 * EPS * fmax(fabs(t), 1e-300))) rc = RC_DTLESSTHANMIN;
    }
    if (active && rc == RC_DEFAULT) {
      bool last = false;
      if (t + dt >= a.t1 - 100.0 * EPS * fabs(a.t1)) { dt = a.t1 - t; last = true; }
      double EEst;
      bool badpiv_step = false;
      if (ROSEN) {
        // ---- Rodas5P step: W = M/(gamma dt) - J factored once, 8 stage solves ------------
        const double cW = 1.0 / (dt * T::GAMMA);
        M::factor(u, p, t, tb, cW, LU);
        badpiv_step = !(M::pivmax(LU) <= MB_PIVOT_GUARD);      // statically ordered LU hit a numerically zero pivot
        njac++;
        if (M::HAS_DT) M::dT(u, p, t, tb, dT);
        const double inv_dt = 1.0 / dt;
#pragma unroll(UNR ? 8 : 1)
        for (int s = 0; s < NST; s++) {
          double sc[N];
#pragma unroll
          for (int i = 0; i < N; i++) { us[i] = u[i]; sc[i] = 0.0; }
#pragma unroll(UNR ? 8 : 1)
          for (int j = 0; j < s; j++) {
            const double aj = T::A(s, j), cj = T::C(s, j);
#pragma unroll
            for (int i = 0; i < N; i++) {
              const double kj = k[j][i];
              us[i] = fma(aj, kj, us[i]);
              sc[i] = fma(cj, kj, sc[i]);
            }
          }
          double r[N];
          M::rhs(us, p, t + T::CT(s) * dt, tb, r);
          const double dd = dt * T::D(s);
#pragma unroll
          for (int i = 0; i < N; i++) {
            double v = r[i];
            if (M::HAS_DT) v = fma(dd, dT[M::HAS_DT ? i : 0], v);
            if (M::isdiff(i)) v = fma(sc[i], inv_dt, v);
            r[i] = v;
          }
          M::solve(LU, cW, p, r);
#pragma unroll
          for (int i = 0; i < N; i++) k[s][i] = r[i];
        }
        nf += NST; nsolve += NST;
#pragma unroll
        for (int i = 0; i < N; i++) un[i] = us[i] + k[NK - 1][i];
        EEst = scaled_rms<N>(k[NK - 1], u, un, a.abstol, a.reltol);
      } else {
        // ---- Tsit5 step (FSAL: k[0] holds f(u,t); refreshed from k[6] on acceptance) -----
#pragma unroll(UNR ? 7 : 1)
        for (int s = 1; s < 7; s++) {
          double acc[N];
#pragma unroll
          for (int i = 0; i < N; i++) acc[i] = 0.0;
#pragma unroll(UNR ? 7 : 1)
          for (int j = 0; j < s; j++) {
            const double aj = cTS_A[s][j];
#pragma unroll
            for (int i = 0; i < N; i++) acc[i] = fma(aj, k[j][i], acc[i]);
          }
#pragma unroll
          for (int i = 0; i < N; i++) us[i] = fma(dt, acc[i], u[i]);
          double r[N];
          M::rhs(us, p, t + cTS_C[s] * dt, tb, r);
#pragma unroll
          for (int i = 0; i < N; i++) k[s][i] = r[i];
        }
        nf += 6;
        double err[N];
#pragma unroll
        for (int i = 0; i < N; i++) {
          double e = 0.0;
#pragma unroll
          for (int j = 0; j < 7; j++) e = fma(TS_BT[j], k[j][i], e);
          err[i] = dt * e;
          un[i] = us[i];
        }
        EEst = scaled_rms<N>(err, u, un, a.abstol, a.reltol);
      }
      if (badpiv_step) {
        rc = RC_UNSTABLE;
      } else if (!isfinite(EEst)) {
        rc = RC_DTNAN;
      } else {
        // ---- PI controller --------------------------------------
        double q, q11 = 0.0;
        if (EEst == 0.0) q = 1.0 / QMAX;
        else {
          q11 = mb_ctl_pow(EEst, BETA1, a.pow_mode);
          q = q11 / mb_ctl_pow(qold, BETA2, a.pow_mode);
          q = fmax(1.0 / QMAX, fmin(1.0 / QMIN, q / SAFETY));
        }
        if (EEst <= 1.0) {
          naccept++;
          const double tn = last ? a.t1 : t + dt;
          while (is < a.n_save && a.saveat[is] <= tn) {       // saveat via dense output
            const double ts = a.saveat[is];
            double* __restrict__ o = a.out_u + (size_t)is * N * a.N + traj;
            if (ts == tn) {
#pragma unroll
              for (int i = 0; i < N; i++) o[(size_t)i * a.N] = un[i];
            } else {
              const double th = (ts - t) / dt;
              if (ROSEN) {
                for (int i = 0; i < N; i++) {
                  double K1 = 0, K2 = 0, K3 = 0;
#pragma unroll
                  for (int j = 0; j < NST; j++) {
                    const double kj = k[j][i];
                    K1 = fma(T::H(0, j), kj, K1); K2 = fma(T::H(1, j), kj, K2); K3 = fma(T::H(2, j), kj, K3);
                  }
                  o[(size_t)i * a.N] = (1.0 - th) * u[i] + th * (un[i] + (1.0 - th) * (K1 + th * (K2 + th * K3)));
                }
              } else {
                double b[7];
#pragma unroll
                for (int j = 0; j < 7; j++)
                  b[j] = th * (TS_R[j][0] + th * (TS_R[j][1] + th * (TS_R[j][2] + th * TS_R[j][3])));
                for (int i = 0; i < N; i++) {
                  double s_ = 0.0;
#pragma unroll
                  for (int j = 0; j < 7; j++) s_ = fma(b[j], k[j][i], s_);
                  o[(size_t)i * a.N] = fma(dt, s_, u[i]);
                }
              }
            }
            is++;
          }
#pragma unroll
          for (int i = 0; i < N; i++) u[i] = un[i];
          if (!ROSEN) {
#pragma unroll
            for (int i = 0; i < N; i++) k[0][i] = k[NK - 1][i];
          }
          t = tn;
          qold = fmax(EEst, QOLDINIT);
          if (a.qs_min <= q && q <= a.qs_max) q = 1.0;      // dead band of step_accept_controller!
          dt = fmin(dt / q, dtmax);
          if (!(t < a.t1)) rc = RC_SUCCESS;
        } else {
          nreject++;
          dt = dt / fmin(1.0 / QMIN, q11 / SAFETY);
        }
      }
    }
    if (active && rc != RC_DEFAULT) {
      // ---- retire this trajectory -----------------------------------------------------------
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

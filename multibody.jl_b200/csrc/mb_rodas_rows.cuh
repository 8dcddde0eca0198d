// Rosenbrock (Rodas5P / Rodas4) ensemble kernel, "row workspace" variant.
//
// Same algorithm and arithmetic as the Rodas5P branch of mbk::ensemble_kernel
// (mb_ensemble.cuh); what differs is where the per-trajectory vectors live.  Every vector
// (8 stage vectors k, stage state, C-sum, u, df/dt, parameters, LU slots) is a row of ONE
// thread-local array that is also indexed dynamically (the rolled stage loop), so the
// compiler keeps all of it in local memory and loads each value where it is used.  In the
// first version the stage state, the C-sum and the right-hand side output (3 x 36 doubles =
// 216 of the 255 registers) were live in registers across the inlined rhs(), which left its
// 2 400-flop body ~40 registers: 300 LDL + 155 STL of spill code, executed 8 times per step.
//
// MB_ROWS_SMEM (set by the model's translation unit) moves rows from local memory to shared memory, laid out
// [component][thread] (conflict-free):
//   1  the C-sum `sc` and df/dt, the two rows written once and read back per stage or per step
//   2  + the stage state `us`; the model then provides rhs_s(), a copy of the right-hand side that reads its state with
//      the [component][thread] stride (MB_RHS_STRIDED)
//   3  only the C-sum          4  mode 1 + the LU slots          5  mode 1, LU slots in a global scratch (a.scratch)
//   6  mode 2 + the state `u` (factor() / dT() read it through MB_FU_AT): 4 N doubles per thread, 221 KB for the robot's
//      192-thread CTA -- the shipped robot configuration (numbers: csrc/mb_model_robot6dof.cu)
#pragma once
#include "mb_ensemble.cuh"
#ifndef MB_ROWS_SMEM
#define MB_ROWS_SMEM 0
#endif
#if !defined(MB_ROWS_NO_CARRY) && !defined(MB_ROWS_CARRY)
#define MB_ROWS_CARRY        // carried stage accumulators (below); MB_ROWS_NO_CARRY selects the first formulation
#endif
#ifndef MB_ROWS_P_SMEM
#define MB_ROWS_P_SMEM 0     // mode 6: number of parameters kept in shared memory (0 or NP; the model's MB_FP_AT must agree)
#endif
#ifndef MB_ROWS_LU_SMEM
#define MB_ROWS_LU_SMEM 0
#endif
#ifndef MB_ROWS_JSPLIT
#define MB_ROWS_JSPLIT 1     // slices of the component range in the stage-combination loop (must divide N)
#endif
#ifndef MB_ROWS_JUNROLL
#define MB_ROWS_JUNROLL 1    // rows of k in flight per slice
#endif

namespace mbk {

template <class M, class T, int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) rosenbrock_rows_kernel(const MbArgs a) {
  constexpr int N = M::N, NST = T::NST;
  constexpr double BETA1 = 7.0 / (10.0 * T::ORDER), BETA2 = 2.0 / (5.0 * T::ORDER);
  constexpr int NPR = (M::NP + N - 1) / N > 0 ? (M::NP + N - 1) / N : 1;
  constexpr bool SLU = MB_ROWS_SMEM == 4;   // mode 4: LU slots in shared memory
  constexpr bool GLU = MB_ROWS_SMEM == 5;   // mode 5: LU slots in a global scratch ([slot][thread] per CTA) kept in the L2 set-aside
  constexpr int NLUR = (SLU || GLU) ? 0 : ((M::NLU + N - 1) / N > 0 ? (M::NLU + N - 1) / N : 1);
  constexpr int R_US = NST, R_SC = NST + 1, R_U = NST + 2, R_DT = NST + 3, R_P = NST + 4, R_LU = R_P + NPR, ROWS = R_LU + NLUR;
  constexpr double EPS = 2.220446049250313e-16;
  constexpr int NH = (N % MB_ROWS_JSPLIT == 0) ? MB_ROWS_JSPLIT : 1, HN = N / NH, JUN = MB_ROWS_JUNROLL;

  double W[ROWS][N];
  double* const us = &W[R_US][0];
  double* const sc = &W[R_SC][0];
  double* const u = &W[R_U][0];
  double* const dT = &W[R_DT][0];
  double* const p = &W[R_P][0];
  constexpr bool SU = MB_ROWS_SMEM == 6;     // mode 6: mode 2 + the state row u (factor / dT read it through MB_FU_AT)
  constexpr bool SROWS = MB_ROWS_SMEM != 0, SUS = MB_ROWS_SMEM == 2 || SU;
  constexpr bool SDT = SROWS && MB_ROWS_SMEM != 3 && M::HAS_DT;            // mode 3: only the C-sum row
  // shared rows: sc | df/dt (its structurally non-zero entries only) | us (modes 2, 6) | u (mode 6) | LU slots (mode 4: all;
  // otherwise the first MB_ROWS_LU_SMEM, addressed by the model's MB_LU_AT)
  constexpr unsigned long long DTM = M::DT_MASK & (N >= 64 ? ~0ull : ((1ull << (N & 63)) - 1ull));
  constexpr int NDTS = SDT ? mb_popc64(DTM) : 0;
  constexpr int S_DT = N, S_US = N + NDTS, S_U = S_US + (SUS ? N : 0), S_LUS = S_U + (SU ? N : 0);
  // mode 6 with MB_ROWS_P_SMEM = NP: the parameters too (read by every right-hand side, factor, dT and solve through MB_FP_AT)
  constexpr bool SP = SU && MB_ROWS_P_SMEM > 0;
  constexpr int S_P = S_LUS + MB_ROWS_LU_SMEM;
  static_assert(!SP || (MB_ROWS_P_SMEM == M::NP && M::IS_ODE), "MB_ROWS_P_SMEM: NP slots, ODE models only (dae_init passes thread-local parameters)");
#ifdef MB_ROWS_LUS0
  static_assert(MB_ROWS_LU_SMEM == 0 || S_LUS == MB_ROWS_LUS0, "MB_LU_AT of the model and the row layout disagree");
#endif
#define MB_DTNZ(i) (((DTM >> ((i) & 63)) & 1ull) != 0ull)
#define MB_DTS(i) (S_DT + mb_popc64(DTM & ((1ull << ((i) & 63)) - 1ull)))
#define MB_US(i) (*(SUS ? &srow[(size_t)(S_US + (i)) * BLOCK] : &us[i]))
#define MB_UU(i) (*(SU ? &srow[(size_t)(S_U + (i)) * BLOCK] : &u[i]))
  double* const srow = mb_dyn_smem + (size_t)M::SLOTS * BLOCK + threadIdx.x;   // [row][component][thread]
  // modes 4 / 5: slot k of this thread is LU[k * BLOCK] (MB_LU_AT of the model's translation unit)
  double* const LU = SLU ? srow + (size_t)S_LUS * BLOCK
                   : GLU ? a.scratch + (size_t)blockIdx.x * ((size_t)M::NLU * BLOCK) + threadIdx.x
                         : &W[0][0] + (size_t)R_LU * N;
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
        if (!M::IS_ODE) {
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
        else {
          dt = initial_dt<M>(*reinterpret_cast<const double(*)[N]>(u), p, tb, a.t0, dtmax, a.abstol, a.reltol, T::ORDER);
          nf += 2;
        }
        if (SU) {
#pragma unroll
          for (int i = 0; i < N; i++) MB_UU(i) = u[i];
        }
        if (SP) {
#pragma unroll
          for (int i = 0; i < M::NP; i++) srow[(size_t)(S_P + i) * BLOCK] = p[i];
        }
        active = true;
        if (!(t < a.t1)) rc = RC_SUCCESS;
        if (init_fail) rc = RC_INITIALFAILURE;
      }
    }
    if (!__syncthreads_or(active)) break;     // block-lockstep stepping (see mb_ensemble.cuh)
    if (active && rc == RC_DEFAULT) {
      ++iter;
      if (iter > maxiters) rc = RC_MAXITERS;
      else if (!isfinite(dt)) rc = RC_DTNAN;
      else if (dt < fmax(a.dtmin, 16.0 * EPS * fmax(fabs(t), 1e-300))) rc = RC_DTLESSTHANMIN;
    }
    if (active && rc == RC_DEFAULT) {
      bool last = false;
      if (t + dt >= a.t1 - 100.0 * EPS * fabs(a.t1)) { dt = a.t1 - t; last = true; }
      const double cW = 1.0 / (dt * T::GAMMA);
      const double* const ust = SU ? srow + (size_t)S_U * BLOCK : u;      // the state as factor() / dT() address it
      const double* const pst = SP ? srow + (size_t)S_P * BLOCK : p;      // the parameters as rhs_s() / factor() / dT() / solve() address them
      M::factor(ust, pst, t, tb, cW, LU);
      const bool badpiv = !(M::pivmax(LU) <= MB_PIVOT_GUARD);     // statically ordered LU hit a numerically zero pivot
      njac++;
      if (M::HAS_DT) {
        M::dT(ust, pst, t, tb, dT);
        if (SDT) {
#pragma unroll
          for (int i = 0; i < N; i++) if (MB_DTNZ(i)) srow[(size_t)MB_DTS(i) * BLOCK] = dT[i];
        }
      }
      const double inv_dt = 1.0 / dt;
#ifdef MB_ROWS_CARRY
      // The accumulators of the NEXT stage are carried across the loop, so the
      // newest row k_s enters them from registers and the loop over stored rows stops at k_{s-2} (7 of the 28 row reads of a
      // Rodas5P step gone); the right-hand side hands its output over in registers; the last stage accumulates the error
      // norm from registers and leaves u_new = us + k_NST in the stage-state row.
      double EEst;
      {
        double x[N], y[N];
        double sacc = 0.0;
#pragma unroll
        for (int i = 0; i < N; i++) { x[i] = MB_UU(i); y[i] = 0.0; }
#pragma unroll 1
        for (int s = 0; s < NST; s++) {
#pragma unroll 1
          for (int j = 0; j + 1 < s; j++) {
            const double aj = T::A(s, j), cj = T::C(s, j);
#pragma unroll
            for (int i = 0; i < N; i++) {
              const double kj = W[j][i];
              x[i] = fma(aj, kj, x[i]);
              y[i] = fma(cj, kj, y[i]);
            }
          }
#pragma unroll
          for (int i = 0; i < N; i++) {
            MB_US(i) = x[i];
            if (SROWS) srow[(size_t)i * BLOCK] = y[i]; else sc[i] = y[i];
          }
          double r[N];
          if (SUS) M::rhs_s(srow + (size_t)S_US * BLOCK, pst, t + T::CT(s) * dt, tb, r);
          else M::rhs(us, p, t + T::CT(s) * dt, tb, r);
          const double dd = dt * T::D(s);
#pragma unroll
          for (int i = 0; i < N; i++) {
            double v = r[i];
            if (M::HAS_DT && MB_DTNZ(i)) v = fma(dd, SDT ? srow[(size_t)MB_DTS(i) * BLOCK] : dT[i], v);
            if (M::isdiff(i)) v = fma(SROWS ? srow[(size_t)i * BLOCK] : sc[i], inv_dt, v);
            r[i] = v;
          }
          M::solve(LU, cW, SP ? pst : p, r);
#pragma unroll
          for (int i = 0; i < N; i++) W[s][i] = r[i];
          {   // unconditional (the last stage writes values nobody reads): x, y must be dead across the right-hand side
            const int sn = s + 1 < NST ? s + 1 : s;
            const double a1 = T::A(sn, s), c1 = T::C(sn, s);
#pragma unroll
            for (int i = 0; i < N; i++) { x[i] = fma(a1, r[i], MB_UU(i)); y[i] = c1 * r[i]; }
          }
          if (s + 1 == NST) {
#pragma unroll
            for (int i = 0; i < N; i++) {
              const double k8 = r[i], ua = MB_UU(i), ub = MB_US(i) + k8;
              MB_US(i) = ub;
              const double scl = a.abstol + fmax(fabs(ua), fabs(ub)) * a.reltol;
              const double q = k8 / scl;
              sacc += q * q;
            }
          }
        }
        EEst = sqrt(sacc / N);
      }
      nf += NST; nsolve += NST;
#define MB_UN(i) MB_US(i)
#else
#pragma unroll 1
      for (int s = 0; s < NST; s++) {
        // stage state us = u + sum_j A[s][j] k_j and C-sum sc = sum_j C[s][j] k_j.  The components are processed in
        // MB_ROWS_JSPLIT slices so that MB_ROWS_JUNROLL rows of k can be in flight at once (registers: 2 slice
        // accumulators + JUNROLL slices of a row): fewer exposed round trips to thread-local memory per stage
#pragma unroll 1
        for (int h = 0; h < NH; h++) {
          const int i0 = h * HN;
          double x[HN], y[HN];
#pragma unroll
          for (int i = 0; i < HN; i++) { x[i] = MB_UU(i0 + i); y[i] = 0.0; }
#pragma unroll(JUN)
          for (int j = 0; j < s; j++) {
            const double aj = T::A(s, j), cj = T::C(s, j);
#pragma unroll
            for (int i = 0; i < HN; i++) {
              const double kj = W[j][i0 + i];
              x[i] = fma(aj, kj, x[i]);
              y[i] = fma(cj, kj, y[i]);
            }
          }
#pragma unroll
          for (int i = 0; i < HN; i++) {
            MB_US(i0 + i) = x[i];
            if (SROWS) srow[(size_t)(i0 + i) * BLOCK] = y[i]; else sc[i0 + i] = y[i];
          }
        }
#ifdef MB_ROWS_RHS_REG
        // variant: the right-hand side hands its output over in registers instead of through the row k[s]
        {
          double r[N];
          if (SUS) M::rhs_s(srow + (size_t)S_US * BLOCK, pst, t + T::CT(s) * dt, tb, r);
          else M::rhs(us, p, t + T::CT(s) * dt, tb, r);
          const double dd = dt * T::D(s);
#pragma unroll
          for (int i = 0; i < N; i++) {
            double v = r[i];
            if (M::HAS_DT && MB_DTNZ(i)) v = fma(dd, SDT ? srow[(size_t)MB_DTS(i) * BLOCK] : dT[i], v);
            if (M::isdiff(i)) v = fma(SROWS ? srow[(size_t)i * BLOCK] : sc[i], inv_dt, v);
            r[i] = v;
          }
#else
        if (SUS) M::rhs_s(srow + (size_t)S_US * BLOCK, pst, t + T::CT(s) * dt, tb, &W[s][0]);
        else M::rhs(us, p, t + T::CT(s) * dt, tb, &W[s][0]);     // du lands in the row that becomes k[s]
        {
          double r[N];
          const double dd = dt * T::D(s);
#pragma unroll
          for (int i = 0; i < N; i++) {
            double v = W[s][i];
            if (M::HAS_DT && MB_DTNZ(i)) v = fma(dd, SDT ? srow[(size_t)MB_DTS(i) * BLOCK] : dT[i], v);
            if (M::isdiff(i)) v = fma(SROWS ? srow[(size_t)i * BLOCK] : sc[i], inv_dt, v);
            r[i] = v;
          }
#endif
          M::solve(LU, cW, SP ? pst : p, r);
#pragma unroll
          for (int i = 0; i < N; i++) W[s][i] = r[i];
        }
      }
      nf += NST; nsolve += NST;
      // u_new = us + k8; error estimate = k8
      double EEst;
      {
        double sacc = 0.0;
#pragma unroll
        for (int i = 0; i < N; i++) {
          const double k8 = W[NST - 1][i], ua = MB_UU(i), ub = MB_US(i) + k8;
          const double scl = a.abstol + fmax(fabs(ua), fabs(ub)) * a.reltol;
          const double q = k8 / scl;
          sacc += q * q;
        }
        EEst = sqrt(sacc / N);
      }
#define MB_UN(i) (MB_US(i) + W[NST - 1][i])
#endif
      if (badpiv) {
        rc = RC_UNSTABLE;
      } else if (!isfinite(EEst)) {
        rc = RC_DTNAN;
      } else {
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
              for (int i = 0; i < N; i++) o[(size_t)i * a.N] = MB_UN(i);
            } else {
              const double th = (ts - t) / dt;
              for (int i = 0; i < N; i++) {
                double K1 = 0, K2 = 0, K3 = 0;
#pragma unroll
                for (int j = 0; j < NST; j++) {
                  const double kj = W[j][i];
                  K1 = fma(T::H(0, j), kj, K1); K2 = fma(T::H(1, j), kj, K2); K3 = fma(T::H(2, j), kj, K3);
                }
                const double un = MB_UN(i);
                o[(size_t)i * a.N] = (1.0 - th) * MB_UU(i) + th * (un + (1.0 - th) * (K1 + th * (K2 + th * K3)));
              }
            }
            is++;
          }
#pragma unroll
          for (int i = 0; i < N; i++) MB_UU(i) = MB_UN(i);
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

#undef MB_US
#undef MB_DTNZ
#undef MB_DTS
#undef MB_UU
#undef MB_UN

}  // namespace mbk

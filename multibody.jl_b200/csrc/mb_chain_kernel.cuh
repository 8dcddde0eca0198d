// Warp-per-trajectory Rosenbrock kernel for long ropes (BASELINE.json config 5: "rope with N = 20-50 links (stiff,
// large state), warp-per-trajectory Rodas5P"; north star (2): "one trajectory per thread (or per warp for large
// chains/ropes)").
//
// One warp integrates one trajectory of n = 126 ... 406 states:
//   * right-hand side: the O(N) articulated-body sweep of mb_chain.h (loops over the links, no generated code);
//   * Jacobian: forward differences, as the reference's rope examples do (`autodiff=false`,
//     docs/src/examples/ropes_and_cables.md:36,61) -- the 32 lanes evaluate 32 perturbed right-hand sides side by
//     side (same instruction stream, different perturbed state: no divergence), ceil((n + 1) / 32) rounds per step;
//   * W = M/(gamma h) - J: dense, column-major in the warp's scratch (global memory, L1/L2 resident: 127 KB at
//     n = 126, 1.3 MB at n = 406), LU with partial pivoting done by the warp -- lanes stride over the rows of a column,
//     so every access of the elimination is a coalesced 256-byte segment;
//   * stage vectors, error norm, dense-output `saveat`: lanes stride over the state components.
// Same step control, tableau and arithmetic order as mb_rodas_rows.cuh / the oracle (Appendix C of SURVEY.md).
// Warps pull trajectories from the global work counter; finished trajectories are reported chunk-wise for the
// streamed read-back like everywhere else.
#pragma once
#include "mb_ensemble.cuh"
#define MBC_FN static __device__ __forceinline__
#include "mb_chain.h"
#include "mb_chain_models.h"

// -DMBC_PROFILE: cycle counters per phase of a warp-step, printed by the warp that retires trajectory 0 (tuning only)
#ifdef MBC_PROFILE
#include <cstdio>
#define MBC_T0() long long t_ph = clock64()
#define MBC_TICK(k) do { const long long t_now = clock64(); prof[k] += t_now - t_ph; t_ph = t_now; } while (0)
#else
#define MBC_T0()
#define MBC_TICK(k)
#endif

namespace mbk {

template <int NLINKS, int ELASTIC>
struct ChainDims {
  static constexpr int N = ELASTIC ? 8 * NLINKS + 6 : 6 * (NLINKS + 1);
  static constexpr int NP = ELASTIC ? 5 : 3;
  static constexpr int NJ = ELASTIC ? 2 * NLINKS + 1 : NLINKS + 1;
  // per-warp scratch, in doubles (every block 32-double aligned)
  static constexpr size_t W = (size_t)N * N;                       // W / LU, column-major
  static constexpr size_t NV = ((size_t)N + 31) & ~(size_t)31;     // one state-sized vector
  static constexpr size_t VECS = 8 + 5;                            // k[8] | u | us | f0 | r | sc
  static constexpr size_t ABA = (size_t)32 * MBC_LINK_DOUBLES * NJ; // sweep scratch of the 32 lanes, lane-interleaved
  static constexpr size_t ABA1 = (size_t)MBC_LINK_DOUBLES * NJ;      // sweep scratch of ONE evaluation (shared memory, stride 1)
  static constexpr size_t PIV = NV;                                // pivot rows (int, in a double-sized slot)
  // The system is second order: half of the rows of W = M/(gamma h) - J are just c x_q - x_v = b_q.  Eliminating them
  // exactly leaves S = c I - J_vv - J_vq / c of HALF the size (8x fewer LU flops, half-length substitution chains);
  // G = J_vq is kept for the right-hand sides of the solves.
  static constexpr int M = N / 2;
  static constexpr size_t SM = ((size_t)M * M + 31) & ~(size_t)31;
  static constexpr int MPAD = (M + 31) & ~31;
  static constexpr size_t PER_WARP = ((W + 31) & ~(size_t)31) + 2 * SM + VECS * NV + ABA + PIV + 32;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// S <- LU(S) with partial pivoting (first maximum, as the oracle's lu_factor); false on a zero / non-finite pivot.
// Column-major, lanes stride over the rows of a column (coalesced).  The trailing update runs over eight columns at a
// time so that eight independent loads are in flight per lane (one column at a time measured one memory round trip per
// fused multiply-add, profiles/r02_rope_long.md).
template <int N>
__device__ bool warp_lu(double* __restrict__ Wm, int* __restrict__ piv, int lane) {
  for (int k = 0; k < N; k++) {
    double* colk = Wm + (size_t)k * N;
    double mx = -1.0;
    int pk = k;
    for (int i = k + lane; i < N; i += 32) {
      const double v = fabs(colk[i]);
      if (v > mx) { mx = v; pk = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double mo = __shfl_xor_sync(0xffffffffu, mx, o);
      const int po = __shfl_xor_sync(0xffffffffu, pk, o);
      if (mo > mx || (mo == mx && po < pk)) { mx = mo; pk = po; }
    }
    if (lane == 0) piv[k] = pk;
    if (!(mx > 0.0) || !isfinite(mx)) return false;
    if (pk != k)
      for (int j = lane; j < N; j += 32) {
        double* c = Wm + (size_t)j * N;
        const double t = c[k]; c[k] = c[pk]; c[pk] = t;
      }
    __syncwarp();
    const double inv = 1.0 / colk[k];
    for (int i = k + 1 + lane; i < N; i += 32) colk[i] *= inv;
    __syncwarp();
    int j = k + 1;
    for (; j + 7 < N; j += 8) {
      double* c0 = Wm + (size_t)j * N;
      double uk[8];
#pragma unroll
      for (int q = 0; q < 8; q++) uk[q] = c0[(size_t)q * N + k];
      for (int i = k + 1 + lane; i < N; i += 32) {
        const double l = colk[i];
        double v[8];
#pragma unroll
        for (int q = 0; q < 8; q++) v[q] = c0[(size_t)q * N + i];
#pragma unroll
        for (int q = 0; q < 8; q++) c0[(size_t)q * N + i] = fma(-l, uk[q], v[q]);
      }
    }
    for (; j < N; j++) {
      double* cj = Wm + (size_t)j * N;
      const double ukj = cj[k];
      for (int i = k + 1 + lane; i < N; i += 32) cj[i] = fma(-colk[i], ukj, cj[i]);
    }
    __syncwarp();
  }
  return true;
}

// y <- S^-1 y with the factors above.  The substitution is a chain of 2 N dependent steps, so everything on the chain
// is on chip: y and the reciprocal diagonal live in shared memory, and the column a step needs was loaded into registers
// during the previous step (a global load inside the step measured ~900 cycles per step, profiles/r02_rope_long.md).
template <int N>
__device__ void warp_lu_solve(const double* __restrict__ Wm, const int* __restrict__ piv, const double* dinv, double* y, int lane) {
  constexpr int R = (N + 31) / 32;
  if (lane == 0)
    for (int k = 0; k < N; k++) {
      const int pk = piv[k];
      if (pk != k) { const double t = y[k]; y[k] = y[pk]; y[pk] = t; }
    }
  double cur[R], nxt[R];
#pragma unroll
  for (int q = 0; q < R; q++) { const int i = lane + 32 * q; cur[q] = (i < N) ? Wm[i] : 0.0; }
  __syncwarp();
  for (int k = 0; k < N; k++) {
    if (k + 1 < N) {
      const double* cn = Wm + (size_t)(k + 1) * N;
#pragma unroll
      for (int q = 0; q < R; q++) { const int i = lane + 32 * q; nxt[q] = (i < N) ? cn[i] : 0.0; }
    }
    const double bk = y[k];
#pragma unroll
    for (int q = 0; q < R; q++) { const int i = lane + 32 * q; if (i > k && i < N) y[i] = fma(-cur[q], bk, y[i]); }
    __syncwarp();
#pragma unroll
    for (int q = 0; q < R; q++) cur[q] = nxt[q];
  }
  {
    const double* c0 = Wm + (size_t)(N - 1) * N;
#pragma unroll
    for (int q = 0; q < R; q++) { const int i = lane + 32 * q; cur[q] = (i < N) ? c0[i] : 0.0; }
  }
  for (int k = N - 1; k >= 0; k--) {
    if (k > 0) {
      const double* cn = Wm + (size_t)(k - 1) * N;
#pragma unroll
      for (int q = 0; q < R; q++) { const int i = lane + 32 * q; nxt[q] = (i < N) ? cn[i] : 0.0; }
    }
    const double bk = y[k] * dinv[k];
#pragma unroll
    for (int q = 0; q < R; q++) { const int i = lane + 32 * q; if (i < k) y[i] = fma(-cur[q], bk, y[i]); }
    __syncwarp();
    if (lane == 0) y[k] = bk;
#pragma unroll
    for (int q = 0; q < R; q++) cur[q] = nxt[q];
  }
  __syncwarp();
}

// Degrees of freedom of the rope: dof d <-> state indices of its coordinate and of its velocity
template <int NLINKS, int ELASTIC>
__device__ __forceinline__ int chain_sq(int d) {
  if (!ELASTIC) return 6 * (d / 3) + d % 3;
  const int i = d >> 2, k = d & 3;
  return (k < 3) ? 8 * i + k : 8 * i + 6;
}
template <int NLINKS, int ELASTIC>
__device__ __forceinline__ int chain_sv(int d) {
  if (!ELASTIC) return 6 * (d / 3) + 3 + d % 3;
  const int i = d >> 2, k = d & 3;
  return (k < 3) ? 8 * i + 3 + k : 8 * i + 7;
}

template <int NLINKS, int ELASTIC, class T, int BLOCK>
#ifndef MBC_MINB
#define MBC_MINB 1      // tuning: minimum resident CTAs per SM the register allocation must allow
#endif
__global__ void __launch_bounds__(BLOCK, MBC_MINB) chain_warp_kernel(const MbArgs a) {
  using Dm = ChainDims<NLINKS, ELASTIC>;
  constexpr int N = Dm::N, NP = Dm::NP, NST = T::NST, M = Dm::M;
  __shared__ double sh_bq[BLOCK / 32][Dm::MPAD], sh_y[BLOCK / 32][Dm::MPAD], sh_di[BLOCK / 32][Dm::MPAD];
  extern __shared__ double mbc_dyn_smem[];      // per warp: sweep scratch of the single-lane (stage) right-hand sides
  double* const bq = sh_bq[threadIdx.x >> 5];
  double* const yv = sh_y[threadIdx.x >> 5];
  double* const dinv = sh_di[threadIdx.x >> 5];
  double* const aba1 = mbc_dyn_smem + (size_t)(threadIdx.x >> 5) * Dm::ABA1;
  constexpr double BETA1 = 7.0 / (10.0 * T::ORDER), BETA2 = 2.0 / (5.0 * T::ORDER);
  constexpr double EPS = 2.220446049250313e-16;
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * BLOCK + threadIdx.x) >> 5;
  double* const base = a.scratch + (size_t)warp * Dm::PER_WARP;
  double* const Wm = base;                                        // raw difference columns f(u + e_j), column-major n x n
  double* const Sm = base + ((Dm::W + 31) & ~(size_t)31);          // Schur complement / its LU, column-major M x M
  double* const Gm = Sm + Dm::SM;                                  // J_vq, column-major M x M
  double* const vec = Gm + Dm::SM;
  double* const K = vec;                       // k[s] = K + s * NV
  double* const u = vec + 8 * Dm::NV;
  double* const us = u + Dm::NV;
  double* const f0 = us + Dm::NV;
  double* const r = f0 + Dm::NV;
  double* const sc = r + Dm::NV;
  double* const aba = sc + Dm::NV;             // lane-interleaved sweep scratch
  int* const piv = reinterpret_cast<int*>(aba + Dm::ABA);
  double* const pp = aba + Dm::ABA + Dm::PIV;   // parameters of this trajectory
  MbcRope R;
  mbc_rope_init(&R, NLINKS, ELASTIC);
  const double dtmax = a.dtmax > 0.0 ? a.dtmax : (a.t1 - a.t0);
  const long long maxiters = a.maxiters > 0 ? a.maxiters : 100000;

  for (;;) {
    unsigned long long idx = 0;
    if (lane == 0) idx = atomicAdd(a.queue, 1ull);
    idx = __shfl_sync(0xffffffffu, idx, 0);
    if (idx >= (unsigned long long)a.N) break;
    const long long traj = (long long)idx;
    for (int i = lane; i < N; i += 32) u[i] = a.u0[(size_t)i * a.N + traj];
    if (lane < NP) pp[lane] = a.p[(size_t)lane * a.N + traj];
    __syncwarp();
    double t = a.t0, dt, qold = QOLDINIT;
    long long iter = 0;
    int is = 0, naccept = 0, nreject = 0, nf = 0, njac = 0, nsolve = 0, rc = RC_DEFAULT;
#ifdef MBC_PROFILE
    long long prof[6] = {0, 0, 0, 0, 0, 0};    // fd columns | W build | LU | stage sums | stage rhs | stage solves
#endif
    while (is < a.n_save && a.saveat[is] <= t) {
      if (a.saveat[is] == t)
        for (int i = lane; i < N; i += 32) a.out_u[((size_t)is * N + i) * a.N + traj] = u[i];
      is++;
    }
    if (a.dt0 > 0.0) dt = a.dt0;
    else {   // Hairer's initial step (same formula as mbk::initial_dt; every state is differential here)
      if (lane == 0) mbc_rope_rhs(&R, u, -1, 0.0, pp, f0, aba, 32);
      __syncwarp();
      double d0 = 0, d1 = 0;
      for (int i = lane; i < N; i += 32) {
        const double sk = a.abstol + fabs(u[i]) * a.reltol;
        d0 += (u[i] / sk) * (u[i] / sk);
        d1 += (f0[i] / sk) * (f0[i] / sk);
      }
      d0 = sqrt(warp_sum(d0) / N); d1 = sqrt(warp_sum(d1) / N);
      double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
      h0 = fmin(h0, dtmax);
      for (int i = lane; i < N; i += 32) us[i] = u[i] + h0 * f0[i];
      __syncwarp();
      if (lane == 0) mbc_rope_rhs(&R, us, -1, 0.0, pp, r, aba, 32);
      __syncwarp();
      double d2 = 0;
      for (int i = lane; i < N; i += 32) {
        const double sk = a.abstol + fabs(u[i]) * a.reltol, q = (r[i] - f0[i]) / sk;
        d2 += q * q;
      }
      d2 = sqrt(warp_sum(d2) / N) / h0;
      const double dm = fmax(d1, d2);
      const double h1 = (dm <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(10.0, -(2.0 + log10(dm)) / (double)T::ORDER);
      dt = fmin(fmin(100.0 * h0, h1), dtmax);
      nf += 2;
    }
    if (!(t < a.t1)) rc = RC_SUCCESS;
    while (rc == RC_DEFAULT) {
      ++iter;
      if (iter > maxiters) { rc = RC_MAXITERS; break; }
      if (!isfinite(dt)) { rc = RC_DTNAN; break; }
      if (dt < fmax(a.dtmin, 16.0 * EPS * fmax(fabs(t), 1e-300))) { rc = RC_DTLESSTHANMIN; break; }
      bool last = false;
      if (t + dt >= a.t1 - 100.0 * EPS * fabs(a.t1)) { dt = a.t1 - t; last = true; }
      const double cW = 1.0 / (dt * T::GAMMA), inv_dt = 1.0 / dt;
      // ---- f0 and the forward-difference Jacobian: lane L of round q evaluates column 32 q + L - 1 (column -1 = f0)
      const double rel = 1.4901161193847656e-08;          // sqrt(eps), as the oracle / FiniteDiff
      MBC_T0();
      for (int q0 = -1; q0 < N; q0 += 32) {
        const int j = q0 + lane;
        if (j < N) {
          const double ej = (j >= 0) ? fmax(rel * fabs(u[j]), rel) : 0.0;
          mbc_rope_rhs(&R, u, j, ej, pp, (j >= 0) ? Wm + (size_t)j * N : f0, aba + lane, 32);
        }
        __syncwarp();
      }
      nf += N + 1; njac++;
      MBC_TICK(0);
      // S = c I - J_vv - J_vq / c and G = J_vq from the difference columns (the rows q' = v are known exactly)
      for (int b = 0; b < M; b++) {
        const int cq = chain_sq<NLINKS, ELASTIC>(b), cv = chain_sv<NLINKS, ELASTIC>(b);
        const double eq = fmax(rel * fabs(u[cq]), rel), ev = fmax(rel * fabs(u[cv]), rel);
        const double* colq = Wm + (size_t)cq * N;
        const double* colv = Wm + (size_t)cv * N;
        for (int a_ = lane; a_ < M; a_ += 32) {
          const int ra = chain_sv<NLINKS, ELASTIC>(a_);
          const double jvq = (colq[ra] - f0[ra]) / eq, jvv = (colv[ra] - f0[ra]) / ev;
          Gm[(size_t)b * M + a_] = jvq;
          Sm[(size_t)b * M + a_] = ((a_ == b) ? cW : 0.0) - jvv - jvq * (dt * T::GAMMA);
        }
      }
      __syncwarp();
      MBC_TICK(1);
      if (!warp_lu<M>(Sm, piv, lane)) { rc = RC_UNSTABLE; break; }
      for (int a_ = lane; a_ < M; a_ += 32) dinv[a_] = 1.0 / Sm[(size_t)a_ * M + a_];
      __syncwarp();
      MBC_TICK(2);
      // ---- stages
      for (int s = 0; s < NST; s++) {
        double* ks = K + (size_t)s * Dm::NV;
        for (int i = lane; i < N; i += 32) {
          double x = u[i], y = 0.0;
          for (int jj = 0; jj < s; jj++) {
            const double kj = K[(size_t)jj * Dm::NV + i];
            x = fma(T::A(s, jj), kj, x);
            y = fma(T::C(s, jj), kj, y);
          }
          us[i] = x; sc[i] = y;
        }
        __syncwarp();
        MBC_TICK(3);
        if (s > 0) {
          if (lane == 0) mbc_rope_rhs(&R, us, -1, 0.0, pp, ks, aba1, 1);
          __syncwarp();
        }
        MBC_TICK(4);
        for (int i = lane; i < N; i += 32) r[i] = fma(sc[i], inv_dt, (s > 0) ? ks[i] : f0[i]);   // autonomous: no df/dt term
        __syncwarp();
        // reduced solve: S x_v = b_v + G b_q / c,  x_q = (b_q + x_v) / c
        for (int a_ = lane; a_ < M; a_ += 32) bq[a_] = r[chain_sq<NLINKS, ELASTIC>(a_)];
        __syncwarp();
        for (int a_ = lane; a_ < M; a_ += 32) {
          double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
          int b = 0;
          for (; b + 3 < M; b += 4) {
            acc0 = fma(Gm[(size_t)b * M + a_], bq[b], acc0);
            acc1 = fma(Gm[(size_t)(b + 1) * M + a_], bq[b + 1], acc1);
            acc2 = fma(Gm[(size_t)(b + 2) * M + a_], bq[b + 2], acc2);
            acc3 = fma(Gm[(size_t)(b + 3) * M + a_], bq[b + 3], acc3);
          }
          for (; b < M; b++) acc0 = fma(Gm[(size_t)b * M + a_], bq[b], acc0);
          yv[a_] = fma((acc0 + acc1) + (acc2 + acc3), dt * T::GAMMA, r[chain_sv<NLINKS, ELASTIC>(a_)]);
        }
        __syncwarp();
        warp_lu_solve<M>(Sm, piv, dinv, yv, lane);
        for (int a_ = lane; a_ < M; a_ += 32) {
          const double xv = yv[a_];
          ks[chain_sq<NLINKS, ELASTIC>(a_)] = (bq[a_] + xv) * (dt * T::GAMMA);
          ks[chain_sv<NLINKS, ELASTIC>(a_)] = xv;
        }
        __syncwarp();
        MBC_TICK(5);
      }
      nf += NST - 1; nsolve += NST;
      const double* k8 = K + (size_t)(NST - 1) * Dm::NV;
      double sacc = 0.0;
      for (int i = lane; i < N; i += 32) {
        const double kk = k8[i], ub = us[i] + kk;
        const double scl = a.abstol + fmax(fabs(u[i]), fabs(ub)) * a.reltol, q = kk / scl;
        sacc += q * q;
      }
      const double EEst = sqrt(warp_sum(sacc) / N);
      if (!isfinite(EEst)) { rc = RC_DTNAN; break; }
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
        while (is < a.n_save && a.saveat[is] <= tn) {        // saveat via dense output
          const double ts = a.saveat[is];
          double* __restrict__ o = a.out_u + (size_t)is * N * a.N + traj;
          const double th = (ts - t) / dt;
          for (int i = lane; i < N; i += 32) {
            const double un = us[i] + k8[i];
            if (ts == tn) o[(size_t)i * a.N] = un;
            else {
              double K1 = 0, K2 = 0, K3 = 0;
              for (int jj = 0; jj < NST; jj++) {
                const double kj = K[(size_t)jj * Dm::NV + i];
                K1 = fma(T::H(0, jj), kj, K1); K2 = fma(T::H(1, jj), kj, K2); K3 = fma(T::H(2, jj), kj, K3);
              }
              o[(size_t)i * a.N] = (1.0 - th) * u[i] + th * (un + (1.0 - th) * (K1 + th * (K2 + th * K3)));
            }
          }
          is++;
        }
        for (int i = lane; i < N; i += 32) u[i] = us[i] + k8[i];
        __syncwarp();
        t = tn;
        qold = fmax(EEst, QOLDINIT);
        if (a.qs_min <= q && q <= a.qs_max) q = 1.0;
        dt = fmin(dt / q, dtmax);
        if (!(t < a.t1)) rc = RC_SUCCESS;
      } else {
        nreject++;
        dt = dt / fmin(1.0 / QMIN, q11 / SAFETY);
      }
    }
    // ---- retire
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    for (; is < a.n_save; is++)
      for (int i = lane; i < N; i += 32) a.out_u[((size_t)is * N + i) * a.N + traj] = qnan;
    __syncwarp();
    if (lane == 0) {
      a.out_tfinal[traj] = t;
      a.retcode[traj] = rc;
      a.stats[0 * a.N + traj] = naccept; a.stats[1 * a.N + traj] = nreject; a.stats[2 * a.N + traj] = nf;
      a.stats[3 * a.N + traj] = njac; a.stats[4 * a.N + traj] = nsolve;
    }
    __threadfence();
    __syncwarp();
    if (lane == 0) mb_retire(a, traj);
#ifdef MBC_PROFILE
    if (lane == 0 && traj == 0) {
      const double st = (double)(naccept + nreject);
      printf("MBC_PROFILE n=%d steps=%d cycles/step: fd %.0f | build %.0f | lu %.0f | sums %.0f | stage rhs %.0f | solves %.0f\n", N,
             naccept + nreject, prof[0] / st, prof[1] / st, prof[2] / st, prof[3] / st, prof[4] / st, prof[5] / st);
    }
#endif
  }
}

}  // namespace mbk

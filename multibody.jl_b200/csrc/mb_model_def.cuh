// MB_DEFINE_MODEL(X, BLOCK, MINB, UNR): wraps the generated functions of model X
// (generated/mbm_X.h) into the traits struct the ensemble kernels expect, instantiates the
// Tsit5 and Rodas5P kernels and registers a launcher.
#pragma once
#include "mb_ensemble.cuh"
#include "mb_rodas_rows.cuh"
#include "mb_registry.h"
#include <cstdlib>

// tuning knob: MBB200_CTAS_PER_SM caps the resident CTAs per SM (persistent grid size)
static inline int mb_cap_ctas_per_sm(int bps) {
  const char* e = getenv("MBB200_CTAS_PER_SM");
  const int cap = e ? atoi(e) : 0;
  return (cap > 0 && cap < bps) ? cap : bps;
}

// MB_RODAS_CLASSIC selects the first Rodas5P kernel (vectors as separate local arrays)
#ifdef MB_RODAS_CLASSIC
#define MB_RODAS_KERNEL(MODEL, BLOCK, MINB, UNR) mbk::ensemble_kernel<MODEL, 2, BLOCK, MINB, UNR>
#else
#define MB_RODAS_KERNEL(MODEL, BLOCK, MINB, UNR) mbk::rodas5p_rows_kernel<MODEL, BLOCK, MINB>
#endif

#define MB_DEFINE_MODEL(X, BLOCK, MINB, UNR)                                                        \
  struct MbModel_##X {                                                                              \
    static constexpr int N = MBM_##X##_N, NP = MBM_##X##_M, NLU = MBM_##X##_NLU,                    \
                         HAS_DT = MBM_##X##_HAS_DT;                                                 \
    static __device__ __forceinline__ void rhs(const double* u, const double* p, double t,         \
                                               const double* tb, double* du) {                      \
      mbm_##X##_rhs(u, p, t, tb, du);                                                               \
    }                                                                                               \
    static __device__ __forceinline__ void dT(const double* u, const double* p, double t,          \
                                              const double* tb, double* o) {                        \
      mbm_##X##_dT(u, p, t, tb, o);                                                                 \
    }                                                                                               \
    static __device__ __forceinline__ void factor(const double* u, const double* p, double t,      \
                                                  const double* tb, double c, double* LU) {         \
      mbm_##X##_factor(u, p, t, tb, c, LU);                                                         \
    }                                                                                               \
    static __device__ __forceinline__ void solve(const double* LU, double c, const double* p,       \
                                                 double* b) {                                       \
      mbm_##X##_solve(LU, c, p, b);                                                                 \
    }                                                                                               \
    static __device__ __forceinline__ int isdiff(int i) { return mbm_##X##_isdiff(i); }             \
  };                                                                                                \
  static cudaError_t mb_launch_##X(int alg, const MbArgs& a, int sm_count, cudaStream_t st) {       \
    int bps = 1;                                                                                    \
    if (alg == 1) {                                                                                 \
      auto kern = mbk::ensemble_kernel<MbModel_##X, 1, BLOCK, MINB, UNR>;                           \
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 0);                \
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, BLOCK, 0);                          \
      bps = mb_cap_ctas_per_sm(bps);                                                                \
      long long grid = (long long)sm_count * (bps > 0 ? bps : 1);                                   \
      long long need = (a.N + BLOCK - 1) / BLOCK;                                                   \
      if (grid > need) grid = need;                                                                 \
      kern<<<(unsigned)grid, BLOCK, 0, st>>>(a);                                                    \
    } else if (alg == 2) {                                                                          \
      auto kern = MB_RODAS_KERNEL(MbModel_##X, BLOCK, MINB, UNR);                                   \
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 0);                \
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, BLOCK, 0);                          \
      bps = mb_cap_ctas_per_sm(bps);                                                                \
      long long grid = (long long)sm_count * (bps > 0 ? bps : 1);                                   \
      long long need = (a.N + BLOCK - 1) / BLOCK;                                                   \
      if (grid > need) grid = need;                                                                 \
      kern<<<(unsigned)grid, BLOCK, 0, st>>>(a);                                                    \
    } else {                                                                                        \
      return cudaErrorInvalidValue;                                                                 \
    }                                                                                               \
    return cudaGetLastError();                                                                      \
  }                                                                                                 \
  static MbKernelEntry mb_entry_##X = {#X, MBM_##X##_N, MBM_##X##_M, MBM_##X##_NLU,                 \
                                       MBM_##X##_IS_ODE,                                            \
                                       {MBM_##X##_F_RHS, MBM_##X##_F_JAC, MBM_##X##_F_DT,           \
                                        MBM_##X##_F_FACTOR, MBM_##X##_F_SOLVE},                     \
                                       mb_launch_##X, nullptr};                                     \
  namespace {                                                                                       \
  struct MbReg_##X {                                                                                \
    MbReg_##X() { mb_register_model(&mb_entry_##X); }                                               \
  } mb_reg_instance_##X;                                                                            \
  }

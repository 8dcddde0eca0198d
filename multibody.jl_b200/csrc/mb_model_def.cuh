// MB_DEFINE_MODEL(X, BLOCK, MINB, UNR): wraps the generated functions of model X
// (generated/mbm_X.h) into the traits struct the ensemble kernels expect, instantiates the
// Tsit5 and Rodas5P kernels and registers a launcher.
#pragma once
#include "mb_ensemble.cuh"
#include "mb_rodas_rows.cuh"
#include "mb_bdf.cuh"
#include "mb_registry.h"
#include <cstdlib>

// tuning knob: MBB200_CTAS_PER_SM caps the resident CTAs per SM (persistent grid size)
static inline int mb_cap_ctas_per_sm(int bps) {
  const char* e = getenv("MBB200_CTAS_PER_SM");
  const int cap = e ? atoi(e) : 0;
  return (cap > 0 && cap < bps) ? cap : bps;
}

// Threads per CTA of the Rosenbrock row kernels.  A lane integrates one trajectory at a time, so a shard of n
// trajectories per CTA takes ceil(n / BLOCK) rounds and the last one runs with idle lanes; the same number of rounds with
// ceil(n / rounds) lanes each keeps every lane busy and leaves the SM less crowded (the per-GPU shard of the 8-GPU strong
// scaling run, 131 072 trajectories = 3.46 rounds of 37 888 lanes: 77.3k trajectories/s with 224 threads against 69.2k
// with 256).  The shared rows keep their BLOCK stride.  MBB200_THREADS_PER_CTA overrides (a multiple of 32).
static inline int mb_threads_per_cta(int block, long long N, long long grid) {
  const char* e = getenv("MBB200_THREADS_PER_CTA");
  const int v = e ? atoi(e) : 0;
  if (v >= 32 && v <= block && v % 32 == 0) return v;
  if (N <= 0 || grid <= 0) return block;
  const long long n = (N + grid - 1) / grid;
  const long long rounds = (n + block - 1) / block;
  long long t = ((n + rounds - 1) / rounds + 31) / 32 * 32;
  if (t < 32) t = 32;
  if (t > block) t = block;
  return (int)t;
}

// Observed variables of saved states: one thread per (save point, trajectory); HBM-bound
// (reads n + m doubles, writes nobs doubles per point; coalesced along the trajectory index).
template <class M, int NOBS>
__global__ void __launch_bounds__(256) mb_observed_kernel(const MbObsArgs a) {
  constexpr int N = M::N, NP = (M::NP > 0 ? M::NP : 1), NO = (NOBS > 0 ? NOBS : 1);
  const long long total = a.N * (long long)a.n_save;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const int is = (int)(g / a.N);
    const long long j = g - (long long)is * a.N;
    double u[N], p[NP], o[NO];
#pragma unroll
    for (int i = 0; i < N; i++) u[i] = a.u[((size_t)is * N + i) * a.N + j];
#pragma unroll
    for (int i = 0; i < M::NP; i++) p[i] = a.p[(size_t)i * a.N + j];
    M::observed(u, p, a.saveat[is], a.tb, o);
#pragma unroll
    for (int k = 0; k < NOBS; k++) a.out[((size_t)is * NOBS + k) * a.N + j] = o[k];
  }
}

// MB_RODAS_CLASSIC selects the first Rodas5P kernel (vectors as separate local arrays)
#ifdef MB_RODAS_CLASSIC
#define MB_RODAS_KERNEL(MODEL, ALG, BLOCK, MINB, UNR) mbk::ensemble_kernel<MODEL, ALG, BLOCK, MINB, UNR>
#else
#define MB_RODAS_KERNEL(MODEL, ALG, BLOCK, MINB, UNR) \
  mbk::rosenbrock_rows_kernel<MODEL, typename mbk::TabOf<ALG>::type, BLOCK, MINB>
#endif

#ifndef MB_ROWS_LU_SMEM
#define MB_ROWS_LU_SMEM 0      // LU slots 0 .. MB_ROWS_LU_SMEM-1 live in shared memory (the model's MB_LU_AT must agree)
#endif
#if defined(MB_RODAS_CLASSIC) || !MB_ROWS_SMEM
#define MB_ROWS_SMEM_BYTES_M(N, HASDT, BLOCK, MASK) 0
#else
#if MB_ROWS_SMEM == 4
#define MB_ROWS_SMEM_BYTES_M(N, HASDT, BLOCK, MASK) (((N) + ((HASDT) ? mb_popc64(MASK) : 0) + MB_ROWS_NLU) * sizeof(double) * (BLOCK))
#else
// rows: C-sum (N) | df/dt (its structurally non-zero entries, not in mode 3) | stage state (modes 2, 6) | state (mode 6) | LU slots (MB_ROWS_LU_SMEM)
#define MB_ROWS_SMEM_BYTES_M(N, HASDT, BLOCK, MASK) (((1 + (MB_ROWS_SMEM == 2 ? 1 : 0) + (MB_ROWS_SMEM == 6 ? 2 : 0)) * (N) + (((HASDT) && MB_ROWS_SMEM != 3) ? mb_popc64(MASK) : 0) + MB_ROWS_LU_SMEM + MB_ROWS_P_SMEM) * sizeof(double) * (BLOCK))
#endif
#endif
// rhs_s: right-hand side reading a [component][thread]-strided state (generated under MB_RHS_STRIDED)
#ifdef MB_RHS_STRIDED
#define MB_RHS_S_NAME(X) mbm_##X##_rhs_s
#else
#define MB_RHS_S_NAME(X) mbm_##X##_rhs
#endif
// shared-memory carve-out in percent of the 228 KB maximum: what the kernel needs, the rest stays L1
static inline int mb_carveout(size_t smem) { return smem ? (int)((smem * 100 + 233471) / 233472) : 0; }
// alg 4 (FBDF): one more instantiation of the generated functions per model; a translation unit whose compile time
// cannot afford it defines MB_NO_BDF and the launcher then answers cudaErrorInvalidValue for alg 4
#ifdef MB_NO_BDF
#define MB_BDF_LAUNCH(MODEL, BLOCK, MINB) return cudaErrorInvalidValue;
#else
#define MB_BDF_LAUNCH(MODEL, BLOCK, MINB)                                                             \
  auto kern = mbk::bdf_kernel<MODEL, BLOCK, MINB>;                                                    \
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                 \
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, mb_carveout(smem));      \
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, BLOCK, smem);                             \
  bps = mb_cap_ctas_per_sm(bps);                                                                      \
  long long grid = (long long)sm_count * (bps > 0 ? bps : 1);                                         \
  long long need = (a.N + BLOCK - 1) / BLOCK;                                                         \
  if (grid > need) grid = need;                                                                       \
  kern<<<(unsigned)grid, BLOCK, smem, st>>>(a);
#endif

#define MB_DEFINE_MODEL(X, BLOCK, MINB, UNR) MB_DEFINE_MODEL_S(X, BLOCK, MINB, UNR, nullptr)
// SCRATCH: size_t (*)(int alg, long long N, int sm_count) or nullptr -- device scratch the kernels of this model need
#define MB_DEFINE_MODEL_S(X, BLOCK, MINB, UNR, SCRATCH)                                               \
  struct MbModel_##X {                                                                              \
    static constexpr int N = MBM_##X##_N, NP = MBM_##X##_M, NLU = MBM_##X##_NLU,                    \
                         HAS_DT = MBM_##X##_HAS_DT, IS_ODE = MBM_##X##_IS_ODE,                      \
                         SLOTS = MBM_##X##_SMEM_SLOTS;                                              \
    static constexpr unsigned long long DT_MASK = MBM_##X##_DT_MASK;                                \
    static __device__ __forceinline__ void rhs(const double* u, const double* p, double t,         \
                                               const double* tb, double* du) {                      \
      mbm_##X##_rhs(u, p, t, tb, du);                                                               \
    }                                                                                               \
    static __device__ __forceinline__ void rhs_s(const double* u, const double* p, double t,       \
                                                 const double* tb, double* du) {                    \
      MB_RHS_S_NAME(X)(u, p, t, tb, du);                                                            \
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
    static __device__ __forceinline__ double pivmax(const double* LU) { return mbm_##X##_pivmax(LU); } \
    static __device__ __forceinline__ void observed(const double* u, const double* p, double t,    \
                                                    const double* tb, double* o) {                  \
      mbm_##X##_observed(u, p, t, tb, o);                                                           \
    }                                                                                               \
  };                                                                                                \
  static cudaError_t mb_launch_##X(int alg, const MbArgs& a, int sm_count, cudaStream_t st) {       \
    int bps = 1;                                                                                    \
    const size_t smem0 = (size_t)MBM_##X##_SMEM_SLOTS * sizeof(double) * BLOCK;                     \
    size_t smem = smem0;                                                                            \
    if (alg == 1) {                                                                                 \
      auto kern = mbk::ensemble_kernel<MbModel_##X, 1, BLOCK, MINB, UNR>;                           \
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);           \
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, mb_carveout(smem)); \
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, BLOCK, smem);                       \
      bps = mb_cap_ctas_per_sm(bps);                                                                \
      long long grid = (long long)sm_count * (bps > 0 ? bps : 1);                                   \
      long long need = (a.N + BLOCK - 1) / BLOCK;                                                   \
      if (grid > need) grid = need;                                                                 \
      kern<<<(unsigned)grid, BLOCK, smem, st>>>(a);                                                    \
    } else if (alg == 2) {                                                                          \
      smem = smem0 + (size_t)MB_ROWS_SMEM_BYTES_M(MBM_##X##_N, MBM_##X##_HAS_DT, BLOCK, MBM_##X##_DT_MASK);                                \
      auto kern = MB_RODAS_KERNEL(MbModel_##X, 2, BLOCK, MINB, UNR);                                   \
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);           \
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, mb_carveout(smem)); \
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, BLOCK, smem);                       \
      bps = mb_cap_ctas_per_sm(bps);                                                                \
      long long grid = (long long)sm_count * (bps > 0 ? bps : 1);                                   \
      const int bt = mb_threads_per_cta(BLOCK, a.N, grid);                                          \
      long long need = (a.N + bt - 1) / bt;                                                         \
      if (grid > need) grid = need;                                                                 \
      kern<<<(unsigned)grid, bt, smem, st>>>(a);                                                    \
    } else if (alg == 3) {                                                                          \
      smem = smem0 + (size_t)MB_ROWS_SMEM_BYTES_M(MBM_##X##_N, MBM_##X##_HAS_DT, BLOCK, MBM_##X##_DT_MASK);                                \
      auto kern = MB_RODAS_KERNEL(MbModel_##X, 3, BLOCK, MINB, UNR);                                   \
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);           \
      cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, mb_carveout(smem)); \
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, BLOCK, smem);                       \
      bps = mb_cap_ctas_per_sm(bps);                                                                \
      long long grid = (long long)sm_count * (bps > 0 ? bps : 1);                                   \
      const int bt = mb_threads_per_cta(BLOCK, a.N, grid);                                          \
      long long need = (a.N + bt - 1) / bt;                                                         \
      if (grid > need) grid = need;                                                                 \
      kern<<<(unsigned)grid, bt, smem, st>>>(a);                                                    \
    } else if (alg == 4) {                                                                          \
      MB_BDF_LAUNCH(MbModel_##X, BLOCK, MINB)                                                       \
    } else {                                                                                        \
      return cudaErrorInvalidValue;                                                                 \
    }                                                                                               \
    return cudaGetLastError();                                                                      \
  }                                                                                                 \
  static cudaError_t mb_launch_obs_##X(const MbObsArgs& a, int sm_count, cudaStream_t st) {         \
    long long blocks = (a.N * (long long)a.n_save + 255) / 256;                                     \
    if (blocks > (long long)sm_count * 8) blocks = (long long)sm_count * 8;                         \
    if (blocks < 1) blocks = 1;                                                                     \
    mb_observed_kernel<MbModel_##X, MBM_##X##_NOBS><<<(unsigned)blocks, 256, 0, st>>>(a);           \
    return cudaGetLastError();                                                                      \
  }                                                                                                 \
  static MbKernelEntry mb_entry_##X = {#X, MBM_##X##_N, MBM_##X##_M, MBM_##X##_NLU,                 \
                                       MBM_##X##_IS_ODE,                                            \
                                       {MBM_##X##_F_RHS, MBM_##X##_F_JAC, MBM_##X##_F_DT,           \
                                        MBM_##X##_F_FACTOR, MBM_##X##_F_SOLVE},                     \
                                       mb_launch_##X, MBM_##X##_NOBS, mb_launch_obs_##X, nullptr, SCRATCH};                                     \
  namespace {                                                                                       \
  struct MbReg_##X {                                                                                \
    MbReg_##X() { mb_register_model(&mb_entry_##X); }                                               \
  } mb_reg_instance_##X;                                                                            \
  }

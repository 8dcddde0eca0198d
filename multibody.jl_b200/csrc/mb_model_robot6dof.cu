// Kernel set of model 'robot6dof' (generated/mbm_robot6dof.h is emitted by codegen/generate.py).
#ifndef MB_ROWS_SMEM
// Rows of the Rosenbrock workspace kept in shared memory, [component][thread] (csrc/mb_rodas_rows.cuh):
//   1  C-sum and df/dt                                  (round 1; 147 KB at 256 threads)
//   6  C-sum, df/dt (its 12 non-zero entries), stage state, state and the 19 parameters: 139 doubles per thread = 214 KB
//      at 192 threads per CTA (shipped).
// Throughput is flat between 192 and 256 threads per SM at equal thread-local traffic and rises as rows leave local
// memory: 606 208 trajectories, same box: mode 1 x 256 threads 84.1k trajectories/s, mode 2 x 224 88.9k, mode 6 x 160
// 89.7k, mode 6 x 192 95.8k (gpurun_out/var_thr2.log -> profiles/r02_robot_rows_modes.log).
#define MB_ROWS_SMEM 6
#endif
#if MB_ROWS_SMEM == 6 && !defined(MB_ROWS_P_SMEM)
#define MB_ROWS_P_SMEM 19   // the 19 swept parameters in shared memory as well (read by every rhs / factor / dT / solve): +1.8 %
#endif
#if MB_ROWS_SMEM == 2 || MB_ROWS_SMEM == 6
#define MB_RHS_STRIDED
#endif
#define MB_NO_BDF   // no FBDF kernel for this model (compile time; the reference uses FBDF on its loop examples only)
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#ifdef MB_R_RHS_NOINLINE
#define MB_FN_RHS static __device__ __noinline__
#endif
#ifdef MB_R_FACTOR_NOINLINE
#define MB_FN_FACTOR static __device__ __noinline__
#endif
#ifdef MB_R_SOLVE_NOINLINE
#define MB_FN_SOLVE static __device__ __noinline__
#endif
#ifndef MB_R_UNR
#define MB_R_UNR 0
#endif
#ifndef MB_R_BLOCK
#define MB_R_BLOCK 192
#endif
#define MB_S_STRIDE MB_R_BLOCK
#if MB_ROWS_SMEM == 4 || MB_ROWS_SMEM == 5
#define MB_LU_AT(k) LU[(k) * MB_S_STRIDE]
#define MB_ROWS_NLU 152
#elif defined(MB_ROWS_LU_SMEM) && MB_ROWS_LU_SMEM > 0
// the first MB_ROWS_LU_SMEM LU slots live in the shared rows behind sc (36) | df/dt (12 non-zero) | us (36) | u (36)
#define MB_ROWS_LUS0 120
#define MB_LU_AT(k) (*((k) < MB_ROWS_LU_SMEM ? (mb_dyn_smem + (size_t)(MB_ROWS_LUS0 + (k)) * MB_S_STRIDE + threadIdx.x) : (LU + (k))))
#endif
#if MB_ROWS_SMEM == 6
#define MB_FU_AT(i) u[(i) * MB_S_STRIDE]
#if defined(MB_ROWS_P_SMEM) && MB_ROWS_P_SMEM > 0
#define MB_FP_AT(i) p[(i) * MB_S_STRIDE]
#endif
#endif
#include "generated/mbm_robot6dof.h"
#ifndef MB_R_MINB
#define MB_R_MINB 1
#endif
#if MB_ROWS_SMEM == 5
// LU slots of the Rosenbrock kernels live in device scratch: [CTA][slot][thread]; one CTA per SM
static size_t mb_robot_scratch(int alg, long long N, int sm_count) {
  if (alg != 2 && alg != 3) return 0;
  long long grid = 1ll * sm_count, need = (N + MB_R_BLOCK - 1) / MB_R_BLOCK;
  if (grid > need) grid = need;
  return (size_t)grid * MB_R_BLOCK * MB_ROWS_NLU * sizeof(double);
}
#define MB_R_SCRATCH mb_robot_scratch
#else
#define MB_R_SCRATCH nullptr
#endif
#define MB_EXPAND(...) MB_DEFINE_MODEL_S(__VA_ARGS__)
MB_EXPAND(robot6dof, MB_R_BLOCK, MB_R_MINB, MB_R_UNR, MB_R_SCRATCH)

// Kernel set of model 'rope6' (C5, documented size: stiff 6-link rope, n = 42).
// Launch geometry measured on the full C5 ensemble (65 536 trajectories, gpurun_out/c5_var.log -> profiles/r02_robot_rows_modes.log):
// no shared rows, 2 x 128 threads per SM 16.4k trajectories/s; rows mode 6 with one 192-thread CTA 16.6k, with one 224-thread
// CTA 18.0k (shipped).  The kernel is dominated by factor() (94.7k of the 224k flops of a step, 58 KB stack frame).
#ifndef MB_V_BLOCK
#define MB_V_BLOCK 224
#endif
#ifndef MB_V_MINB
#define MB_V_MINB 1
#endif
#ifndef MB_ROWS_SMEM
#define MB_ROWS_SMEM 6      // C-sum, stage state, state and the 3 parameters in shared memory (129 doubles per thread, 226 KB)
#endif
#if MB_ROWS_SMEM == 2 || MB_ROWS_SMEM == 6
#define MB_RHS_STRIDED
#endif
#if MB_ROWS_SMEM == 6 && !defined(MB_ROWS_P_SMEM)
#define MB_ROWS_P_SMEM 3
#endif
#define MB_NO_BDF   // no FBDF kernel for this model (compile time; the reference uses FBDF on its loop examples only)
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE MB_V_BLOCK
#if MB_ROWS_SMEM == 6
#define MB_FU_AT(i) u[(i) * MB_S_STRIDE]
#define MB_FP_AT(i) p[(i) * MB_S_STRIDE]
#endif
#include "generated/mbm_rope6.h"
#define MB_EXPAND(...) MB_DEFINE_MODEL(__VA_ARGS__)
MB_EXPAND(rope6, MB_V_BLOCK, MB_V_MINB, 0)

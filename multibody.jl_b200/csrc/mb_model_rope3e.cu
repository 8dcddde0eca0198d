// Kernel set of model 'rope3e' (flexible 3-link rope of test/basic_tests.jl:928-959, n = 30).
#ifdef MB_VAR_CLASSIC
#define MB_RODAS_CLASSIC
#endif
#ifndef MB_VAR_BLOCK
#define MB_VAR_BLOCK 128
#endif
#ifndef MB_VAR_MINB
#define MB_VAR_MINB 1
#endif
#ifndef MB_ROWS_SMEM
#define MB_ROWS_SMEM 0      // 6 = C-sum, stage state, state and the 5 parameters in shared memory (95 doubles per thread): measured
                            // 15.0k (192 threads) / 16.7k (256 threads) against 15.9k for this setting on the full C5E ensemble -- not adopted
#endif
#if MB_ROWS_SMEM == 2 || MB_ROWS_SMEM == 6
#define MB_RHS_STRIDED
#endif
#if MB_ROWS_SMEM == 6 && !defined(MB_ROWS_P_SMEM)
#define MB_ROWS_P_SMEM 5
#endif
#define MB_NO_BDF   // no FBDF kernel for this model (compile time; the reference uses FBDF on its loop examples only)
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE MB_VAR_BLOCK
#if MB_ROWS_SMEM == 6
#define MB_FU_AT(i) u[(i) * MB_S_STRIDE]
#define MB_FP_AT(i) p[(i) * MB_S_STRIDE]
#endif
#include "generated/mbm_rope3e.h"
MB_DEFINE_MODEL(rope3e, MB_VAR_BLOCK, MB_VAR_MINB, 0)

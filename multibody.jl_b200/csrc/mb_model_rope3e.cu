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
#define MB_NO_BDF   // no FBDF kernel for this model (compile time; the reference uses FBDF on its loop examples only)
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE MB_VAR_BLOCK
#include "generated/mbm_rope3e.h"
MB_DEFINE_MODEL(rope3e, MB_VAR_BLOCK, MB_VAR_MINB, 0)

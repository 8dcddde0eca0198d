// Kernel set of model 'rope3e' (flexible 3-link rope of test/basic_tests.jl:928-959, n = 30).
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_rope3e.h"
MB_DEFINE_MODEL(rope3e, 128, 1, 0)

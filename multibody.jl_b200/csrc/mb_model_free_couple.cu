// Kernel set of model 'free_couple' (test/test_worldforces.jl force-couple cases: free BodyShape,
// 12 Euler-angle root states; Tsit5 / Rodas4).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_free_couple.h"
MB_DEFINE_MODEL(free_couple, 128, 2, 1)

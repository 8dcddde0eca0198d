// Kernel set of model 'three_springs' (C2, second model: quaternion root body + massless spring
// node, index-1 DAE; Rodas5P / Rodas4).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_three_springs.h"
MB_DEFINE_MODEL(three_springs, 128, 2, 1)

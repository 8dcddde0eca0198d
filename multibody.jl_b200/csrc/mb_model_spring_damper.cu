// Kernel set of model 'spring_damper' (generated/mbm_spring_damper.h is emitted by codegen/generate.py).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_spring_damper.h"
MB_DEFINE_MODEL(spring_damper, 128, 2, 1)

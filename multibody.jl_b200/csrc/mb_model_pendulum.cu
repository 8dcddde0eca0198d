// Kernel set of model 'pendulum' (generated/mbm_pendulum.h is emitted by codegen/generate.py).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_pendulum.h"
MB_DEFINE_MODEL(pendulum, 128, 4, 1)

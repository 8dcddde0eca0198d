// Kernel set of model 'fourbar' (C3: closed loop, index-1 mass-matrix DAE; Rodas5P only).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_fourbar.h"
MB_DEFINE_MODEL(fourbar, 128, 1, 0)

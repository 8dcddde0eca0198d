// Kernel set of model 'rope6' (C5, documented size: stiff 6-link rope, n = 42).
#define MB_NO_BDF   // no FBDF kernel for this model (compile time; the reference uses FBDF on its loop examples only)
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_rope6.h"
MB_DEFINE_MODEL(rope6, 128, 1, 0)

// Kernel set of model 'furuta' (test/test_orientation_getters.jl:3-47: two revolutes,
// two BodyShapes, tip body, axis dampers; n = 4; Rodas5P).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_furuta.h"
MB_DEFINE_MODEL(furuta, 128, 2, 1)

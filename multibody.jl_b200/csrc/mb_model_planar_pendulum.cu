// Kernel set of model 'planar_pendulum' (test/test_PlanarMechanics.jl:38-61: planar pendulum,
// n = 2; Rodas5P).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_planar_pendulum.h"
MB_DEFINE_MODEL(planar_pendulum, 128, 2, 1)

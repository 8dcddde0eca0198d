// Kernel set of model 'fourbar_planar' (planar loop closed by RevolutePlanarLoopConstraint,
// docs/src/examples/kinematic_loops.md:38-80; index-1 DAE, Rodas5P / Rodas4).
#define MB_RODAS_CLASSIC   // small state: everything fits in registers, the first kernel is faster
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#define MB_S_STRIDE 128
#include "generated/mbm_fourbar_planar.h"
MB_DEFINE_MODEL(fourbar_planar, 128, 1, 0)

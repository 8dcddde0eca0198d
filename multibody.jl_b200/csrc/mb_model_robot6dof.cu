// Kernel set of model 'robot6dof' (generated/mbm_robot6dof.h is emitted by codegen/generate.py).
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#include "generated/mbm_robot6dof.h"
MB_DEFINE_MODEL(robot6dof, 128, 1, 0)

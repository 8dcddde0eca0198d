// Kernel set of model 'robot6dof' (generated/mbm_robot6dof.h is emitted by codegen/generate.py).
#include "mb_model_def.cuh"
#define MB_FN static __device__ __forceinline__
#define MB_RESTRICT __restrict__
#define MB_NO_PATTERN
#include "generated/mbm_robot6dof.h"
#ifndef MB_R_BLOCK
#define MB_R_BLOCK 256
#endif
#ifndef MB_R_MINB
#define MB_R_MINB 1
#endif
#define MB_EXPAND(...) MB_DEFINE_MODEL(__VA_ARGS__)
MB_EXPAND(robot6dof, MB_R_BLOCK, MB_R_MINB, 0)

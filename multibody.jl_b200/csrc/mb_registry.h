// Registry of compiled-in model kernel sets (one entry per generated model).
#pragma once
#include <cuda_runtime.h>
#include "mb_common.cuh"

struct MbKernelEntry {
  const char* name;
  int n, m, nlu, is_ode;
  long long flops[5];   // rhs, jac, dT, factor, solve (algorithmic)
  // launches the ensemble kernel for `alg` on `stream`; grid sized to fill the device
  cudaError_t (*launch)(int alg, const MbArgs& args, int sm_count, cudaStream_t stream);
  int nobs;             // observed (eliminated) variables the generated code can evaluate
  cudaError_t (*launch_observed)(const MbObsArgs& args, int sm_count, cudaStream_t stream);
  const MbKernelEntry* next;
  // device scratch the kernel of `alg` needs for a shard of N trajectories on a device with sm_count SMs
  // (0 / null for the thread-per-trajectory kernels, whose working state is registers + local memory)
  size_t (*scratch_bytes)(int alg, long long N, int sm_count);
};
void mb_register_model(MbKernelEntry* e);
const MbKernelEntry* mb_find_model(const char* name);

// Kernel sets of the long-rope models (mb_chain_models.h): warp-per-trajectory Rodas5P / Rodas4 over the
// loop-based right-hand side of mb_chain.h.  Registered with the runtime like the generated models.
#include <cstdlib>
#include "mb_chain_kernel.cuh"
#include "mb_registry.h"

#define MBC_BLOCK 64      // 2 warps = 2 trajectories per CTA; the grid is persistent (warps pull work)

template <int NLINKS, int ELASTIC>
static size_t chain_smem() { return (size_t)(MBC_BLOCK / 32) * mbk::ChainDims<NLINKS, ELASTIC>::ABA1 * sizeof(double); }
template <int NLINKS, int ELASTIC, class T>
static long long chain_grid(long long N, int sm_count) {
  auto kern = mbk::chain_warp_kernel<NLINKS, ELASTIC, T, MBC_BLOCK>;
  int bps = 1;
  const size_t smem = chain_smem<NLINKS, ELASTIC>();
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, MBC_BLOCK, smem);
  const char* e = getenv("MBB200_CTAS_PER_SM");
  const int cap = e ? atoi(e) : 0;
  if (cap > 0 && cap < bps) bps = cap;
  long long grid = (long long)sm_count * (bps > 0 ? bps : 1);
  // every warp works in ~0.5-3 MB of scratch: more resident warps hide more latency, fewer keep the scratch in the 126 MB L2.
  // MBB200_CHAIN_CTAS overrides the number of CTAs (tuning: profiles/r02_rope_long.md)
  const char* e2 = getenv("MBB200_CHAIN_CTAS");
  const long long want = e2 ? atoll(e2) : 0;
  if (want > 0 && want < grid) grid = want;
  const long long need = (N + (MBC_BLOCK / 32) - 1) / (MBC_BLOCK / 32);
  if (grid > need) grid = need;
  return grid < 1 ? 1 : grid;
}
template <int NLINKS, int ELASTIC>
static size_t chain_scratch(int alg, long long N, int sm_count) {
  const long long g = (alg == 3) ? chain_grid<NLINKS, ELASTIC, mbk::TabRodas4>(N, sm_count)
                                 : chain_grid<NLINKS, ELASTIC, mbk::TabRodas5P>(N, sm_count);
  return (size_t)g * (MBC_BLOCK / 32) * mbk::ChainDims<NLINKS, ELASTIC>::PER_WARP * sizeof(double);
}
template <int NLINKS, int ELASTIC>
static cudaError_t chain_launch(int alg, const MbArgs& a, int sm_count, cudaStream_t st) {
  if (!a.scratch) return cudaErrorInvalidValue;
  if (alg == 2) {
    const long long g = chain_grid<NLINKS, ELASTIC, mbk::TabRodas5P>(a.N, sm_count);
    mbk::chain_warp_kernel<NLINKS, ELASTIC, mbk::TabRodas5P, MBC_BLOCK><<<(unsigned)g, MBC_BLOCK, chain_smem<NLINKS, ELASTIC>(), st>>>(a);
  } else if (alg == 3) {
    const long long g = chain_grid<NLINKS, ELASTIC, mbk::TabRodas4>(a.N, sm_count);
    mbk::chain_warp_kernel<NLINKS, ELASTIC, mbk::TabRodas4, MBC_BLOCK><<<(unsigned)g, MBC_BLOCK, chain_smem<NLINKS, ELASTIC>(), st>>>(a);
  } else {
    return cudaErrorNotSupported;     // ropes are stiff: Rosenbrock methods only
  }
  return cudaGetLastError();
}

// Algorithmic flops (hand count of mb_chain.h, cost table of BASELINE.md section 4): ~1500 per spherical joint
// (three sincos, 3x3 Cholesky, articulated-inertia update and its transform), ~700 per prismatic joint.
// jac = n forward differences; factor = jac + dense LU 2/3 n^3; solve = 2 n^2.
#define MBC_RHS_FLOPS(NL, EL) (1500ll * ((NL) + 1) + ((EL) ? 700ll * (NL) : 0ll))
#define MBC_NST(NL, EL) ((EL) ? 8 * (NL) + 6 : 6 * ((NL) + 1))
#define MBC_REGISTER(NAME, NL, EL)                                                                                     \
  static MbKernelEntry mb_entry_##NAME = {#NAME, MBC_NST(NL, EL), (EL) ? 5 : 3, 0, 1,                                  \
      {MBC_RHS_FLOPS(NL, EL), (long long)MBC_NST(NL, EL) * MBC_RHS_FLOPS(NL, EL), 0,                                   \
       (long long)MBC_NST(NL, EL) * MBC_RHS_FLOPS(NL, EL) + 2ll * MBC_NST(NL, EL) * MBC_NST(NL, EL) * MBC_NST(NL, EL) / 3, \
       2ll * MBC_NST(NL, EL) * MBC_NST(NL, EL)},                                                                       \
      chain_launch<NL, EL>, 0, nullptr, nullptr, chain_scratch<NL, EL>};                                               \
  namespace { struct MbReg_##NAME { MbReg_##NAME() { mb_register_model(&mb_entry_##NAME); } } mb_reg_instance_##NAME; }
MBC_CHAIN_MODELS(MBC_REGISTER)

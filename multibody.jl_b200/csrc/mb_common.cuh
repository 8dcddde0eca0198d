// Shared device helpers: time-table lookup, method coefficients, launch descriptor.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

// ---- shared time tables ---------------------------------------------------------------
// Uniform-grid lookups of the reference's KinematicPTP interpolants
// (/root/reference/src/robot/path_planning.jl:191-193: CubicSpline, LinearInterpolation,
// both with ExtrapolationType.Constant).  The model image carries (value, z) pairs per knot
// (include/mbb200.h); mbb200_model_load converts them to the device layout used here:
//   tb[0..7]  = t0, h, nk, ntab, 0...                       (64-byte header)
//   then (nk + 1) interval rows x ntab tables x 4 coefficients c0..c3 of
//   S(a) = c0 + a (c1 + a (c2 + a c3)),  a = t - t_i        (one 32-byte sector per lookup)
// Row 0 is "before the first knot" and row nk "after the last" (constant extrapolation:
// c1..c3 = 0).  All tables of one row are contiguous, and a right-hand side evaluates all its
// lookups at one time point, so the segment is located once (MB_SEG) and shared.
#define MB_TB_HEADER 8
struct MbSeg {
  const double* row;   // coefficients of this interval, all tables
  double a;            // t - t_i
};
__device__ __forceinline__ MbSeg mb_seg(const double* __restrict__ tb, double t) {
  const double t0 = __ldg(tb), h = __ldg(tb + 1);
  const int nk = (int)__ldg(tb + 2), ntab = (int)__ldg(tb + 3);
  const double x = (t - t0) / h;
  int row, i = (int)x;
  if (i > nk - 2) i = nk - 2;
  if (x <= 0.0) { row = 0; i = 0; }
  else if (x >= (double)(nk - 1)) row = nk;
  else row = i + 1;
  MbSeg sg;
  sg.row = tb + MB_TB_HEADER + (size_t)row * ntab * 4;
  sg.a = (x - i) * h;
  return sg;
}
template <int DERIV>
__device__ __forceinline__ double mb_tab(const MbSeg& sg, int table) {
  const double2* c = reinterpret_cast<const double2*>(sg.row + table * 4);
  const double2 c01 = __ldg(c), c23 = __ldg(c + 1);
  const double a = sg.a;
  if (DERIV == 0) return fma(a, fma(a, fma(a, c23.y, c23.x), c01.y), c01.x);
  if (DERIV == 1) return fma(a, fma(a, 3.0 * c23.y, 2.0 * c23.x), c01.y);
  if (DERIV == 2) return fma(a, 6.0 * c23.y, 2.0 * c23.x);
  if (DERIV == 3) return 6.0 * c23.y;
  return 0.0;
}
#define MB_SEG_T MbSeg
#define MB_SEG(tb, t) mb_seg(tb, t)
#define MB_TAB_CUBIC0(tb, sg, k) mb_tab<0>(sg, k)
#define MB_TAB_CUBIC1(tb, sg, k) mb_tab<1>(sg, k)
#define MB_TAB_CUBIC2(tb, sg, k) mb_tab<2>(sg, k)
#define MB_TAB_CUBIC3(tb, sg, k) mb_tab<3>(sg, k)
#define MB_TAB_LINEAR0(tb, sg, k) mb_tab<0>(sg, k)
#define MB_TAB_LINEAR1(tb, sg, k) mb_tab<1>(sg, k)
#define MB_TAB_LINEAR2(tb, sg, k) mb_tab<2>(sg, k)
#define MB_TAB_LINEAR3(tb, sg, k) mb_tab<3>(sg, k)
#define MB_SINCOS(x, s, c) sincos(x, &(s), &(c))
// staged temporaries of the generated code (codegen/emit.py): slot k of this thread, in dynamic
// shared memory laid out [slot][thread] (conflict-free); MB_S_STRIDE = threads per CTA of the TU
__host__ __device__ constexpr int mb_popc64(unsigned long long m) { int c = 0; for (; m; m &= m - 1) c++; return c; }
extern __shared__ double mb_dyn_smem[];
#define MB_S_DECL
#define MB_S(k) ((volatile double*)mb_dyn_smem)[(k) * MB_S_STRIDE + threadIdx.x]   // volatile: no store-to-load forwarding back into registers

// ---- method coefficients (published methods) ------------------------
// __constant__ tables: indices are warp-uniform, so every read is a constant-bank broadcast.
namespace mbk {

// Tsit5 (Tsitouras 2011)
static __constant__ double TS_BT[7] = {-0.00178001105222577714, -0.0008164344596567469, 0.007880878010261995,
                             -0.1447110071732629, 0.5823571654525552, -0.45808210592918697,
                             0.015151515151515152};
static __constant__ double TS_R[7][4] = {{1.0, -2.763706197274826, 2.9132554618219126, -1.0530884977290216},
                               {0.0, 0.13169999999999998, -0.2234, 0.1017},
                               {0.0, 3.9302962368947516, -5.941033872131505, 2.490627285651253},
                               {0.0, -12.411077166933676, 30.33818863028232, -16.548102889244902},
                               {0.0, 37.50931341651104, -88.1789048947664, 47.37952196281928},
                               {0.0, -27.896526289197286, 65.09189467479366, -34.87065786149661},
                               {0.0, 1.5, -4.0, 2.5}};

// Rodas5P (Steinebach 2023)
constexpr double RD_GAMMA = 0.21193756319429014;
static __constant__ double RD_H[3][8] = {
    {25.948786856663858, -2.5579724845846235, 10.433815404888879, -2.3679251022685204, 0.524948541321073, 1.1241088310450404, 0.4272876194431874, -0.17202221070155493},
    {-9.91568850695171, -0.9689944594115154, 3.0438037242978453, -24.495224566215796, 20.176138334709044, 15.98066361424651, -6.789040303419874, -6.710236069923372},
    {11.419903575922262, 2.8879645146136994, 72.92137995996029, 80.12511834622643, -52.072871366152654, -59.78993625266729, -0.15582684282751913, 4.883087185713722}};

// PI step controller of the reference's integrators
// (beta1 = 7/(10 order), beta2 = 2/(5 order) are set per method inside the kernels)
constexpr double SAFETY = 0.9, QMIN = 0.2, QMAX = 10.0,
                 QOLDINIT = 1e-4;
// largest admissible |1 / pivot| of the statically ordered LU (generated pivmax()): beyond it W is numerically singular
constexpr double MB_PIVOT_GUARD = 1e10;

// `fastpower` of the reference's PI controller (FastPower.jl, formerly DiffEqBase.fastpow):
// Float32 exp2 of y times an approximate log2 (rational fit on the significand, D. Goldberg,
// "Fast approximate logarithms III", table 2 line 8).  Relative error ~1e-5 in the step-size
// factor; restated so that step sequences follow the reference's, not just converged solutions.
__device__ __forceinline__ float mb_fastlog2(float x) {
  const float a = 0.338953f, b = 2.198599f, c = 1.523692f;
  const unsigned ux1i = __float_as_uint(x);
  const unsigned ex = (ux1i & 0x7F800000u) >> 23;
  const bool greater = (ux1i & 0x00400000u) != 0u;
  float signif = __uint_as_float((ux1i & 0x007FFFFFu) | (greater ? 0x3f000000u : 0x3f800000u));
  const float fexp = __fsub_rn((float)ex, greater ? 126.0f : 127.0f);
  signif = __fsub_rn(signif, 1.0f);
  // no FMA contraction: the CPU restatement (oracle, -ffp-contract=off) rounds each operation
  return __fadd_rn(fexp, __fdiv_rn(__fmul_rn(signif, __fadd_rn(__fmul_rn(a, signif), b)), __fadd_rn(signif, c)));
}
__device__ __forceinline__ double mb_ctl_pow(double x, double y, int pow_mode) {
  if (pow_mode == 1) return pow(x, y);
  if (x == 0.0) return 0.0;
  // exp2 in double, rounded to Float32: bit-identical to the CPU restatement (exp2f implementations differ by an ulp)
  return (double)(float)exp2((double)__fmul_rn((float)y, mb_fastlog2((float)x)));
}

enum { RC_DEFAULT = 0, RC_SUCCESS = 1, RC_DTNAN = 3, RC_MAXITERS = 4, RC_DTLESSTHANMIN = 5, RC_UNSTABLE = 6, RC_INITIALFAILURE = 7 };

}  // namespace mbk

// ---- launch descriptor (all device pointers) ---------------------------------------------
struct MbArgs {
  long long N;               // trajectories in this shard
  const double* u0;          // [n][N]
  const double* p;           // [m][N]
  const double* tb;          // packed tables
  const double* saveat;      // [n_save]
  int n_save;
  double* out_u;             // [n_save][n][N]
  double* out_tfinal;        // [N]
  int* retcode;              // [N]
  int* stats;                // [5][N]
  unsigned long long* queue; // work counter (trajectory index dispenser)
  double t0, t1, abstol, reltol, dt0, dtmin, dtmax;
  long long maxiters;
  double qs_min, qs_max;     // dead band of the PI controller (step kept when qs_min <= q <= qs_max)
  int pow_mode;              // 0 fastpower (reference), 1 exact pow
  // streamed read-back (mb_runtime.cu): per-chunk counters of retired trajectories; the thread that
  // completes a chunk of 2^chunk_shift trajectories raises its flag in mapped host memory
  unsigned* done;            // nullable
  int* flags;
  int chunk_shift;
  double* scratch;           // kernel scratch of models that ask for it (MbKernelEntry::scratch_bytes), else null
};

// called by the thread that has just written all results of trajectory `traj`
__device__ __forceinline__ void mb_retire(const MbArgs& a, long long traj) {
  if (!a.done) return;
  __threadfence();                                           // results visible device-wide before the count
  const long long c = traj >> a.chunk_shift;
  long long size = a.N - (c << a.chunk_shift);
  if (size > (1ll << a.chunk_shift)) size = 1ll << a.chunk_shift;
  if ((long long)atomicAdd(&a.done[c], 1u) + 1 == size) {
    __threadfence_system();
    *(volatile int*)&a.flags[c] = 1;
  }
}

// ---- observed-variable evaluation (the step after the solve: `sol(t, idxs = expr)`) ----------
struct MbObsArgs {
  long long N;            // trajectories
  int n_save;
  const double* u;        // [n_save][n][N] saved states
  const double* p;        // [m][N]
  const double* tb;       // packed tables
  const double* saveat;   // [n_save]
  double* out;            // [n_save][nobs][N]
};

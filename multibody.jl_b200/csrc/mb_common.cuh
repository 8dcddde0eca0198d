// Shared device helpers: time-table lookup, method coefficients, launch descriptor.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

// ---- shared time tables ---------------------------------------------------------------
// Uniform-grid lookups of the reference's KinematicPTP interpolants
// (/root/reference/src/robot/path_planning.jl:191-193: CubicSpline, LinearInterpolation,
// both with ExtrapolationType.Constant).  Packed layout: tb[0]=t0, tb[1]=dt, tb[2]=nk,
// then per table nk interleaved (value, z) pairs -- one 32-byte sector per lookup.
template <int CUBIC, int DERIV>
__device__ __forceinline__ double mb_interp(const double* __restrict__ tb, int table, double t) {
  const double t0 = __ldg(tb), h = __ldg(tb + 1);
  const int nk = (int)__ldg(tb + 2);
  const double x = (t - t0) / h;
  const double* base = tb + 3 + (size_t)table * 2 * nk;
  if (x <= 0.0) return DERIV == 0 ? __ldg(base) : 0.0;
  if (x >= (double)(nk - 1)) return DERIV == 0 ? __ldg(base + 2 * (nk - 1)) : 0.0;
  int i = (int)x;
  if (i > nk - 2) i = nk - 2;
  const double a = (x - i) * h, b = h - a;
  const double v0 = __ldg(base + 2 * i), z0 = __ldg(base + 2 * i + 1);
  const double v1 = __ldg(base + 2 * i + 2), z1 = __ldg(base + 2 * i + 3);
  if (!CUBIC) {
    if (DERIV == 0) return (v0 * b + v1 * a) / h;
    return DERIV == 1 ? (v1 - v0) / h : 0.0;
  }
  const double c0 = v0 / h - z0 * h / 6.0, c1 = v1 / h - z1 * h / 6.0;
  if (DERIV == 0) return z0 * b * b * b / (6 * h) + z1 * a * a * a / (6 * h) + c1 * a + c0 * b;
  if (DERIV == 1) return -z0 * b * b / (2 * h) + z1 * a * a / (2 * h) + c1 - c0;
  if (DERIV == 2) return (z0 * b + z1 * a) / h;
  if (DERIV == 3) return (z1 - z0) / h;
  return 0.0;
}
#define mb_interp_cubic0(tb, k, t) mb_interp<1, 0>(tb, k, t)
#define mb_interp_cubic1(tb, k, t) mb_interp<1, 1>(tb, k, t)
#define mb_interp_cubic2(tb, k, t) mb_interp<1, 2>(tb, k, t)
#define mb_interp_cubic3(tb, k, t) mb_interp<1, 3>(tb, k, t)
#define mb_interp_linear0(tb, k, t) mb_interp<0, 0>(tb, k, t)
#define mb_interp_linear1(tb, k, t) mb_interp<0, 1>(tb, k, t)
#define mb_interp_linear2(tb, k, t) mb_interp<0, 2>(tb, k, t)
#define mb_interp_linear3(tb, k, t) mb_interp<0, 3>(tb, k, t)

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
constexpr double BETA1 = 7.0 / 50.0, BETA2 = 2.0 / 25.0, SAFETY = 0.9, QMIN = 0.2, QMAX = 10.0,
                 QOLDINIT = 1e-4;

enum { RC_DEFAULT = 0, RC_SUCCESS = 1, RC_DTNAN = 3, RC_MAXITERS = 4, RC_DTLESSTHANMIN = 5, RC_UNSTABLE = 6 };

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
};

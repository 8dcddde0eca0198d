/* ORACLE -- test infrastructure only.
 * Shared time tables: restates DataInterpolations [un-vendored, >= 8.6.1] `CubicSpline`
 * (natural), `LinearInterpolation`, both with `ExtrapolationType.Constant`, as used at
 * /root/reference/src/robot/path_planning.jl:191-193.
 * Packed layout: tb[0]=t0, tb[1]=dt, tb[2]=nk, then per table nk pairs (value, z)
 * where z is the spline second derivative at the knot (0 for linear tables). */
#ifndef MBO_INTERP_H
#define MBO_INTERP_H
#include <math.h>
#include <stddef.h>
static inline double mbo_interp(const double* tb, int table, double t, int cubic, int deriv) {
  const double t0 = tb[0], h = tb[1];
  const int nk = (int)tb[2];
  const double* T = tb + 3 + (size_t)table * 2 * nk;
  double x = (t - t0) / h;
  if (x <= 0.0) return deriv == 0 ? T[0] : 0.0;
  if (x >= (double)(nk - 1)) return deriv == 0 ? T[2 * (nk - 1)] : 0.0;
  int i = (int)x; if (i > nk - 2) i = nk - 2;
  double a = (x - i) * h, b = h - a;
  double v0 = T[2 * i], z0 = T[2 * i + 1], v1 = T[2 * i + 2], z1 = T[2 * i + 3];
  if (!cubic) {
    if (deriv == 0) return (v0 * b + v1 * a) / h;
    return deriv == 1 ? (v1 - v0) / h : 0.0;
  }
  double c0 = v0 / h - z0 * h / 6.0, c1 = v1 / h - z1 * h / 6.0;
  switch (deriv) {
    case 0: return z0 * b * b * b / (6 * h) + z1 * a * a * a / (6 * h) + c1 * a + c0 * b;
    case 1: return -z0 * b * b / (2 * h) + z1 * a * a / (2 * h) + c1 - c0;
    case 2: return (z0 * b + z1 * a) / h;
    case 3: return (z1 - z0) / h;
    default: return 0.0;
  }
}
#define mb_interp_cubic0(tb, k, t) mbo_interp(tb, k, t, 1, 0)
#define mb_interp_cubic1(tb, k, t) mbo_interp(tb, k, t, 1, 1)
#define mb_interp_cubic2(tb, k, t) mbo_interp(tb, k, t, 1, 2)
#define mb_interp_cubic3(tb, k, t) mbo_interp(tb, k, t, 1, 3)
#define mb_interp_linear0(tb, k, t) mbo_interp(tb, k, t, 0, 0)
#define mb_interp_linear1(tb, k, t) mbo_interp(tb, k, t, 0, 1)
#define mb_interp_linear2(tb, k, t) mbo_interp(tb, k, t, 0, 2)
#define mb_interp_linear3(tb, k, t) mbo_interp(tb, k, t, 0, 3)
/* generated-header hooks: the oracle locates the segment per lookup and evaluates the
 * DataInterpolations formulas above (the product shares the segment and uses the expanded
 * polynomial, csrc/mb_common.cuh) */
#define MB_SEG_T double
#define MB_SEG(tb, t) (t)
#define MB_TAB_CUBIC0(tb, sg, k) mbo_interp(tb, k, sg, 1, 0)
#define MB_TAB_CUBIC1(tb, sg, k) mbo_interp(tb, k, sg, 1, 1)
#define MB_TAB_CUBIC2(tb, sg, k) mbo_interp(tb, k, sg, 1, 2)
#define MB_TAB_CUBIC3(tb, sg, k) mbo_interp(tb, k, sg, 1, 3)
#define MB_TAB_LINEAR0(tb, sg, k) mbo_interp(tb, k, sg, 0, 0)
#define MB_TAB_LINEAR1(tb, sg, k) mbo_interp(tb, k, sg, 0, 1)
#define MB_TAB_LINEAR2(tb, sg, k) mbo_interp(tb, k, sg, 0, 2)
#define MB_TAB_LINEAR3(tb, sg, k) mbo_interp(tb, k, sg, 0, 3)
/* staged temporaries of the generated code: plain locals here (shared memory on the GPU) */
#define MB_S_DECL double mb_s_[128];
#define MB_S(k) mb_s_[k]
#define MB_SINCOS(x, s, c) do { (s) = sin(x); (c) = cos(x); } while (0)
#endif

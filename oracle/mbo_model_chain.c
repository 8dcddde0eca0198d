/* ORACLE -- test infrastructure only. Never linked into the product.
 * Long ropes (BASELINE.json config 5: 20 and 50 links): the loop-based right-hand side of
 * multibody.jl_b200/csrc/mb_chain.h (articulated-body sweep over `Rope`, /root/reference/src/components.jl:574-626)
 * compiled as plain C.  These models have no analytic Jacobian: the integrator differences the right-hand side
 * forward, as the reference's rope examples do (`Rodas4(autodiff=false)`, docs/src/examples/ropes_and_cables.md:36,61).
 * `rope6c` / `rope3ec` are the same code at the sizes of the generated models rope6 / rope3e, which pin it
 * (tests/test_chain.py). */
#include <string.h>
#include "mbo.h"
#include "../multibody.jl_b200/csrc/mb_chain.h"
#include "../multibody.jl_b200/csrc/mb_chain_models.h"

static void chain_rhs(int N, int elastic, const double* u, const double* p, double* du) {
  MbcRope R;
  mbc_rope_init(&R, N, elastic);
  double ws[MBC_LINK_DOUBLES * (2 * MBC_MAX_LINKS + 1)];
  mbc_rope_rhs(&R, u, -1, 0.0, p, du, ws, 1);
}
static const int ONES[8 * MBC_MAX_LINKS + 6] = {[0 ... 8 * MBC_MAX_LINKS + 5] = 1};
static const int NOPAT[1] = {0};
#define CHAIN_MODEL(NAME, N, EL)                                                                                    \
  static void rhs_##NAME(const double* u, const double* p, double t, const double* tb, double* du) {                 \
    (void)t; (void)tb;                                                                                                \
    chain_rhs(N, EL, u, p, du);                                                                                       \
  }                                                                                                                   \
  static void dT_##NAME(const double* u, const double* p, double t, const double* tb, double* o) {                    \
    (void)u; (void)p; (void)t; (void)tb;                                                                              \
    memset(o, 0, sizeof(double) * ((EL) ? 8 * (N) + 6 : 6 * ((N) + 1)));   /* autonomous */                            \
  }                                                                                                                   \
  const mbo_model mbo_model_##NAME = {#NAME, (EL) ? 8 * (N) + 6 : 6 * ((N) + 1), (EL) ? 5 : 3, 0, NOPAT, NOPAT, ONES, \
                                      rhs_##NAME, 0 /* no analytic Jacobian */, dT_##NAME, 0, 0, 0, 0, 0};
MBC_CHAIN_MODELS(CHAIN_MODEL)

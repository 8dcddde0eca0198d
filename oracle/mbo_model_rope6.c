/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'rope6': the generated right-hand side / Jacobian (csrc/generated/mbm_rope6.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_rope6.h"
const mbo_model mbo_model_rope6 = {"rope6", MBM_rope6_N, MBM_rope6_M, MBM_rope6_NNZ, mbm_rope6_jrow, mbm_rope6_jcol, mbm_rope6_mass,
                                 mbm_rope6_rhs, mbm_rope6_jac, mbm_rope6_dT, MBM_rope6_NLU, mbm_rope6_factor, mbm_rope6_solve,
                                 MBM_rope6_NOBS, mbm_rope6_observed, mbm_rope6_pivmax};

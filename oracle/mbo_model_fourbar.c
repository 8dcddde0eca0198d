/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'fourbar': the generated right-hand side / Jacobian (csrc/generated/mbm_fourbar.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_fourbar.h"
const mbo_model mbo_model_fourbar = {"fourbar", MBM_fourbar_N, MBM_fourbar_M, MBM_fourbar_NNZ, mbm_fourbar_jrow, mbm_fourbar_jcol, mbm_fourbar_mass,
                                 mbm_fourbar_rhs, mbm_fourbar_jac, mbm_fourbar_dT, MBM_fourbar_NLU, mbm_fourbar_factor, mbm_fourbar_solve,
                                 MBM_fourbar_NOBS, mbm_fourbar_observed, mbm_fourbar_pivmax};

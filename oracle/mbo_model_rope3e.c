/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'rope3e': the generated right-hand side / Jacobian (csrc/generated/mbm_rope3e.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_rope3e.h"
const mbo_model mbo_model_rope3e = {"rope3e", MBM_rope3e_N, MBM_rope3e_M, MBM_rope3e_NNZ, mbm_rope3e_jrow, mbm_rope3e_jcol, mbm_rope3e_mass,
                                 mbm_rope3e_rhs, mbm_rope3e_jac, mbm_rope3e_dT, MBM_rope3e_NLU, mbm_rope3e_factor, mbm_rope3e_solve,
                                 MBM_rope3e_NOBS, mbm_rope3e_observed, mbm_rope3e_pivmax};

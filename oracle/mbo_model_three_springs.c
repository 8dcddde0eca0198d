/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'three_springs': the generated right-hand side / Jacobian (csrc/generated/mbm_three_springs.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_three_springs.h"
const mbo_model mbo_model_three_springs = {"three_springs", MBM_three_springs_N, MBM_three_springs_M, MBM_three_springs_NNZ, mbm_three_springs_jrow, mbm_three_springs_jcol, mbm_three_springs_mass,
                                 mbm_three_springs_rhs, mbm_three_springs_jac, mbm_three_springs_dT, MBM_three_springs_NLU, mbm_three_springs_factor, mbm_three_springs_solve,
                                 MBM_three_springs_NOBS, mbm_three_springs_observed, mbm_three_springs_pivmax};

/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'spring_damper': the generated right-hand side / Jacobian (csrc/generated/mbm_spring_damper.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_spring_damper.h"
const mbo_model mbo_model_spring_damper = {"spring_damper", MBM_spring_damper_N, MBM_spring_damper_M, MBM_spring_damper_NNZ, mbm_spring_damper_jrow, mbm_spring_damper_jcol, mbm_spring_damper_mass,
                                 mbm_spring_damper_rhs, mbm_spring_damper_jac, mbm_spring_damper_dT, MBM_spring_damper_NLU, mbm_spring_damper_factor, mbm_spring_damper_solve,
                                 MBM_spring_damper_NOBS, mbm_spring_damper_observed, mbm_spring_damper_pivmax};

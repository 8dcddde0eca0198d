/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'fourbar_planar': the generated right-hand side / Jacobian (csrc/generated/mbm_fourbar_planar.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_fourbar_planar.h"
const mbo_model mbo_model_fourbar_planar = {"fourbar_planar", MBM_fourbar_planar_N, MBM_fourbar_planar_M, MBM_fourbar_planar_NNZ, mbm_fourbar_planar_jrow, mbm_fourbar_planar_jcol, mbm_fourbar_planar_mass,
                                 mbm_fourbar_planar_rhs, mbm_fourbar_planar_jac, mbm_fourbar_planar_dT, MBM_fourbar_planar_NLU, mbm_fourbar_planar_factor, mbm_fourbar_planar_solve,
                                 MBM_fourbar_planar_NOBS, mbm_fourbar_planar_observed, mbm_fourbar_planar_pivmax};

/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'furuta': the generated right-hand side / Jacobian (csrc/generated/mbm_furuta.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_furuta.h"
const mbo_model mbo_model_furuta = {"furuta", MBM_furuta_N, MBM_furuta_M, MBM_furuta_NNZ, mbm_furuta_jrow, mbm_furuta_jcol, mbm_furuta_mass,
                                 mbm_furuta_rhs, mbm_furuta_jac, mbm_furuta_dT, MBM_furuta_NLU, mbm_furuta_factor, mbm_furuta_solve,
                                 MBM_furuta_NOBS, mbm_furuta_observed, mbm_furuta_pivmax};

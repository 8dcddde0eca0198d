/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'free_couple': the generated right-hand side / Jacobian (csrc/generated/mbm_free_couple.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_free_couple.h"
const mbo_model mbo_model_free_couple = {"free_couple", MBM_free_couple_N, MBM_free_couple_M, MBM_free_couple_NNZ, mbm_free_couple_jrow, mbm_free_couple_jcol, mbm_free_couple_mass,
                                 mbm_free_couple_rhs, mbm_free_couple_jac, mbm_free_couple_dT, MBM_free_couple_NLU, mbm_free_couple_factor, mbm_free_couple_solve,
                                 MBM_free_couple_NOBS, mbm_free_couple_observed, mbm_free_couple_pivmax};

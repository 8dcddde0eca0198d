/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'robot6dof': the generated right-hand side / Jacobian (csrc/generated/mbm_robot6dof.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_robot6dof.h"
const mbo_model mbo_model_robot6dof = {"robot6dof", MBM_robot6dof_N, MBM_robot6dof_M, MBM_robot6dof_NNZ, mbm_robot6dof_jrow, mbm_robot6dof_jcol, mbm_robot6dof_mass,
                                 mbm_robot6dof_rhs, mbm_robot6dof_jac, mbm_robot6dof_dT, MBM_robot6dof_NLU, mbm_robot6dof_factor, mbm_robot6dof_solve,
                                 MBM_robot6dof_NOBS, mbm_robot6dof_observed, mbm_robot6dof_pivmax};

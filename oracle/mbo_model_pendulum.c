/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'pendulum': the generated right-hand side / Jacobian (csrc/generated/mbm_pendulum.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_pendulum.h"
const mbo_model mbo_model_pendulum = {"pendulum", MBM_pendulum_N, MBM_pendulum_M, MBM_pendulum_NNZ, mbm_pendulum_jrow, mbm_pendulum_jcol, mbm_pendulum_mass,
                                 mbm_pendulum_rhs, mbm_pendulum_jac, mbm_pendulum_dT, MBM_pendulum_NLU, mbm_pendulum_factor, mbm_pendulum_solve,
                                 MBM_pendulum_NOBS, mbm_pendulum_observed, mbm_pendulum_pivmax};

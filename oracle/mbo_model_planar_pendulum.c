/* ORACLE -- test infrastructure only. Never linked into the product.
 * Model 'planar_pendulum': the generated right-hand side / Jacobian (csrc/generated/mbm_planar_pendulum.h, emitted
 * from the restated reference equations) compiled as plain C. */
#include "mbo.h"
#include "mbo_interp.h"
#define MB_FN static
#define MB_RESTRICT restrict
#include "../multibody.jl_b200/csrc/generated/mbm_planar_pendulum.h"
const mbo_model mbo_model_planar_pendulum = {"planar_pendulum", MBM_planar_pendulum_N, MBM_planar_pendulum_M, MBM_planar_pendulum_NNZ, mbm_planar_pendulum_jrow, mbm_planar_pendulum_jcol, mbm_planar_pendulum_mass,
                                 mbm_planar_pendulum_rhs, mbm_planar_pendulum_jac, mbm_planar_pendulum_dT, MBM_planar_pendulum_NLU, mbm_planar_pendulum_factor, mbm_planar_pendulum_solve,
                                 MBM_planar_pendulum_NOBS, mbm_planar_pendulum_observed, mbm_planar_pendulum_pivmax};

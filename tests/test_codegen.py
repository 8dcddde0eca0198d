"""Code generator: DAG differentiation, the emitted Jacobian / time derivative / structured
factor+solve pair (compiled as plain C inside the oracle library), table lookups."""
import numpy as np
import pytest

from multibody_jl_b200.codegen.dag import Graph, Ex, sin, cos, sqrt, atan2, fmax
from multibody_jl_b200.mechanics.zoo import MODELS, get_model


def _point(sysm, rng):
    u = sysm.u0 + 0.3 * rng.normal(size=sysm.n)
    p = sysm.p0 * (1 + 0.2 * rng.uniform(-1, 1, size=sysm.m)) if sysm.m else np.zeros(1)
    return u, p, float(rng.uniform(0.05, 1.9))


def test_dag_diff_against_finite_differences():
    g = Graph()
    x = [Ex(g, g.input("u", i)) for i in range(3)]
    f = [sin(x[0] * x[1]) / (2.0 + cos(x[2])) + sqrt(x[0] * x[0] + 1.0) * x[2],
         atan2(x[1], x[0] + 3.0) - fmax(x[2], 0.1) * x[0]]
    d = g.diff([e.n for e in f], [v.n for v in x])
    pt = np.array([0.7, -0.4, 1.3])

    def ev(nodes, v):
        return np.array(g.evaluate(nodes, {("u", i): float(v[i]) for i in range(3)}, None))
    for i in range(2):
        for j in range(3):
            h = 1e-6
            e = np.zeros(3); e[j] = h
            fd = (ev([f[i].n], pt + e)[0] - ev([f[i].n], pt - e)[0]) / (2 * h)
            an = ev([d[i].get(j, g.zero)], pt)[0]
            assert abs(fd - an) < 1e-8 * max(1.0, abs(an))


def test_hash_consing_and_flop_count():
    g = Graph()
    a, b = Ex(g, g.input("u", 0)), Ex(g, g.input("u", 1))
    assert (a * b + 1.0).n == (a * b + 1.0).n
    assert g.flops([(a * b + a / b).n]) == 1 + 1 + 4       # mul, add, div (BASELINE.md cost table)


@pytest.mark.parametrize("name", sorted(MODELS))
def test_generated_jacobian_and_dT_match_finite_differences(oracle, name):
    sysm = get_model(name)
    tb = sysm.tables.pack()
    rng = np.random.default_rng(1)
    for _ in range(3):
        u, p, t = _point(sysm, rng)
        J = oracle.jac_dense(name, u, p, t, tb)
        f = lambda v, tt=t: oracle.eval_fn(name, 0, v, p, tt, tb)
        for j in range(sysm.n):
            scale = np.maximum(np.abs(J[:, j]), 1.0) * max(1.0, np.abs(J).max() * 1e-6)
            errs = []
            for h0 in (1e-6, 1e-4, 1e-3):       # the larger step for ill-conditioned mass matrices (300 kg on a 1.7 kg link: round-off in f ~ 1e-9)
                h = h0 * max(1.0, abs(u[j]))
                e = np.zeros(sysm.n); e[j] = h
                fd = (f(u + e) - f(u - e)) / (2 * h)
                errs.append((np.abs(fd - J[:, j]) / scale).max())
            assert min(errs) < (1e-4 if sysm.n < 30 else 5e-4), (name, j)   # FD truncation: h^2 f''' grows with 1/inertia (free_couple) and the rope accelerations
        dT = oracle.eval_fn(name, 2, u, p, t, tb)
        h = 1e-7
        fd = (f(u, t + h) - f(u, t - h)) / (2 * h)
        assert np.abs(fd - dT).max() <= 1e-4 * max(1.0, np.abs(dT).max())


@pytest.mark.parametrize("name", sorted(MODELS))
def test_generated_factor_solve_is_the_inverse_of_W(oracle, name):
    """mbm_X_factor/solve (static pivot sequence chosen at code-generation time) must solve
    (c M - J) x = b to working precision along the whole range of step sizes."""
    sysm = get_model(name)
    tb = sysm.tables.pack()
    M = np.diag(oracle.mass_diag(name).astype(float))
    rng = np.random.default_rng(2)
    for h in (1e-6, 1e-4, 1e-2, 0.3):
        u, p, t = _point(sysm, rng)
        c = 1.0 / (0.21193756319429014 * h)
        W = c * M - oracle.jac_dense(name, u, p, t, tb)
        b = rng.normal(size=sysm.n)
        x = oracle.factor_solve(name, u, p, t, tb, c, b)
        want = np.linalg.solve(W, b)
        assert np.abs(x - want).max() <= 1e-9 * max(1.0, np.abs(want).max()) * max(1.0, np.linalg.cond(W) * 1e-7), (name, h)


def test_generated_rhs_equals_dag_interpretation(oracle):
    for name in ("pendulum", "spring_damper"):
        sysm = get_model(name)
        rng = np.random.default_rng(3)
        u, p, t = _point(sysm, rng)
        np.testing.assert_allclose(oracle.eval_fn(name, 0, u, p, t, sysm.tables.pack()), sysm.eval_rhs(u, p, t),
                                   rtol=1e-12, atol=1e-12)


def test_table_lookup_c_equals_python_and_is_a_natural_spline(oracle):
    sysm = get_model("robot6dof")
    T = sysm.tables
    assert T.nk == 401 and abs(T.dt - 0.01) < 1e-15          # src/robot/path_planning.jl:29
    z = T.z[0]
    assert z[0] == 0.0 and z[-1] == 0.0                       # natural end conditions
    for t in (0.005, 0.11742, 0.5, 1.2386, 1.999, 3.99):
        i = int(t / T.dt)
        # knots are interpolated exactly; inside, the cubic is within h^2 max|q''| / 8 of the chord
        assert abs(T.eval("cubic", 0, 0, i * T.dt) - T.values[0][i]) < 1e-12
    # clamp outside [0, 4] (ExtrapolationType.Constant, path_planning.jl:191-193)
    assert T.eval("cubic", 0, 0, -1.0) == T.values[0][0] and T.eval("cubic", 0, 0, 5.0) == T.values[0][-1]
    assert T.eval("linear", 6, 0, 5.0) == T.values[6][-1]


def test_flop_counts_in_header_are_consistent():
    import os
    import re
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(here, "multibody.jl_b200", "csrc", "generated", "mbm_robot6dof.h")).read()
    d = {k: int(v) for k, v in re.findall(r"#define MBM_robot6dof_(\w+) (\d+)", text)}
    assert d["N"] == 36 and d["M"] == 19
    assert d["F_FACTOR"] >= d["F_JAC"] > d["F_RHS"] > 0
    assert d["F_SOLVE"] < 2 * 36 * 36          # structured solve beats the dense 2 n^2


def test_df_dt_mask_and_strided_accessors_in_the_robot_header(oracle):
    """The Rosenbrock row kernel keeps only the structurally non-zero rows of df/dt in shared memory
    (MBM_*_DT_MASK, csrc/mb_rodas_rows.cuh) and reads state / parameters through MB_FU_AT / MB_FP_AT.  The mask must be
    exactly the rows whose df/dt is not identically zero: 2 per axis (the controller rows the reference path enters)."""
    import os
    import re
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(here, "multibody.jl_b200", "csrc", "generated", "mbm_robot6dof.h")).read()
    mask = int(re.search(r"#define MBM_robot6dof_DT_MASK 0x([0-9a-f]+)ull", text).group(1), 16)
    assert bin(mask).count("1") == 12
    body = text[text.index("void mbm_robot6dof_dT("):]
    body = body[:body.index("\n}\n")]
    zero_rows = {int(i) for i in re.findall(r"dT\[(\d+)\] = 0\.0;", body)}
    assert zero_rows == {i for i in range(36) if not (mask >> i) & 1}
    # numerically: finite-difference df/dt of the oracle's right-hand side vanishes exactly on the masked-out rows
    from multibody_jl_b200 import ensembles
    spec = ensembles.make_ensemble("C4", 1)
    sysm = get_model("robot6dof")
    u, p, tb = spec.u0[:, 0].copy(), spec.p[:, 0].copy(), sysm.tables.pack()
    t, h = 0.37, 1e-6                                                        # inside the acceleration phase of the path
    d = (oracle.eval_fn("robot6dof", 0, u, p, t + h, tb) - oracle.eval_fn("robot6dof", 0, u, p, t - h, tb)) / (2 * h)
    assert all(d[i] == 0.0 for i in zero_rows) and all(d[i] != 0.0 for i in range(36) if (mask >> i) & 1)
    for acc in ("MB_FU_AT(", "MB_FP_AT("):
        assert acc in text[text.index("void mbm_robot6dof_factor("):]      # factor / solve read through the accessors
    plain = text[text.index("void mbm_robot6dof_rhs("):text.index("#ifndef MB_FP_AT")]   # the plain rhs (initial step, Tsit5,
    assert "MB_FP_AT(" not in plain and "p[" in plain                                   # observed) keeps thread-local parameters

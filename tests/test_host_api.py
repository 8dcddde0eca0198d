"""Host-side mirror of the reference API (multibody -> ODEProblem -> EnsembleProblem) and the
synthetic ensembles of SURVEY.md 8d.  CPU only; nothing here calls a compute entry."""
import numpy as np
import pytest

from multibody_jl_b200 import api, ensembles
from multibody_jl_b200.mechanics.robot import point_to_point
from multibody_jl_b200.mechanics.zoo import get_model


def test_multibody_and_odeproblem_maps():
    ssys = api.multibody("pendulum")
    assert ssys.state_names == ["joint.phi", "joint.w"]          # docs/src/examples/pendulum.md:53
    prob = api.ODEProblem(ssys, {"joint.phi": 0.3, "damper.d": 10.0}, (0.0, 1.0))
    assert prob.u0[0] == 0.3 and prob.p[ssys.param_index("damper.d")] == 10.0 and prob.tspan == (0.0, 1.0)
    with pytest.raises(KeyError):
        api.ODEProblem(ssys, {"nonexistent": 1.0})
    with pytest.raises(KeyError):
        api.multibody("no_such_model")
    assert api.multibody(ssys) is ssys


def test_state_counts_match_the_reference():
    assert get_model("robot6dof").n == 36          # docs/src/examples/robot.md:135
    assert get_model("spring_damper").n == 14      # 12 (free body, Euler) + 2 (prismatic)
    assert get_model("pendulum").n == 2


def test_ensemble_problem_prob_func_and_matrices():
    ssys = api.multibody("pendulum")
    prob = api.ODEProblem(ssys, tspan=(0.0, 1.0))
    e = api.EnsembleProblem(prob, prob_func=lambda pr, i, rep: api.remake(pr, u0=[0.1 * i, 0.0]))
    U, P = e.materialise(5)
    np.testing.assert_allclose(U[0], 0.1 * np.arange(5))
    assert U.shape == (2, 5) and P.shape[1] == 5
    U2, _ = api.EnsembleProblem(prob, u0_matrix=U).materialise(5)
    np.testing.assert_array_equal(U, U2)


def test_ensemble_draws_are_shard_independent():
    whole = ensembles.make_ensemble("C4", 64)
    a = ensembles.make_ensemble("C4", 40)
    b = ensembles.make_ensemble("C4", 24, offset=40)
    np.testing.assert_array_equal(np.concatenate([a.p, b.p], axis=1), whole.p)
    np.testing.assert_array_equal(np.concatenate([a.u0, b.u0], axis=1), whole.u0)
    sysm = get_model("robot6dof")
    m = whole.p[sysm.param_index("mechanics.load.m")]
    assert 5.0 <= m.min() and m.max() < 25.0 and m.std() > 3.0
    big = ensembles.uniform(np.arange(200000), 3, 0.0, 1.0)
    assert abs(big.mean() - 0.5) < 5e-3 and abs(big.var() - 1 / 12) < 5e-3


def test_config_sizes():
    assert ensembles.make_ensemble("C1", 1).n_total == 1024
    assert ensembles.make_ensemble("C2", 1).n_total == 1 << 20
    assert ensembles.make_ensemble("C4", 1).n_total == 1 << 20
    assert len(ensembles.make_ensemble("C2", 1).saveat) == 100
    assert len(ensembles.make_ensemble("C4", 1).saveat) == 21


def test_point_to_point_limits_and_end_points():
    """Restates test/test_traj.jl:4-46."""
    t = np.arange(-1.0, 3.0, 0.001)
    q1, qd_max, qdd_max = np.array([1.0, 1.2]), np.array([0.7, 1.2]), np.array([0.9, 1.1])
    q, qd, qdd, t1 = point_to_point(t, np.zeros(2), q1, qd_max, qdd_max)
    assert (q <= q1 + 1e-12).all() and (qd <= qd_max + 1e-12).all() and (qdd <= qdd_max + 1e-12).all()
    last_moving = t[np.nonzero(qd.any(axis=1))[0][-1]]
    assert abs(last_moving - t1) < 1e-3 * t1 + 2e-3
    rng = np.random.default_rng(0)
    N = 200
    t0 = rng.normal()
    q0, q1 = rng.normal(size=N), rng.normal(size=N)
    qd_max, qdd_max = rng.random(N) + 1e-3, rng.random(N) + 1e-3
    *_, t1 = point_to_point([0.0], q0, q1, qd_max, qdd_max, t0)
    t = np.arange(t0 - 1, t1 + 0.1, 0.01)
    q, qd, qdd, _ = point_to_point(t, q0, q1, qd_max, qdd_max, t0)
    np.testing.assert_allclose(q[0], q0, atol=1e-3)
    np.testing.assert_allclose(q[-1], q1, atol=1e-3)
    assert not qd[0].any() and not qd[-1].any() and not qdd[0].any() and not qdd[-1].any()
    assert (q >= np.minimum(q0, q1) - 1e-4).all() and (q <= np.maximum(q0, q1) + 1e-4).all()
    assert (np.abs(qd) <= 1.001 * qd_max).all() and (np.abs(qdd) <= 1.001 * qdd_max).all()


def test_robot_reference_path_switch_times():
    """the default motion switches at t = 0.11742, 1.23862, 1.35605 s."""
    from multibody_jl_b200.mechanics import robot as R
    t = np.arange(401) * 0.01
    Q, Qd, Qdd, te = point_to_point(t, np.deg2rad(R.Q0_DEG), np.deg2rad(R.Q1_DEG), R.QD_MAX, R.QDD_MAX)
    assert abs(te - 1.35605) < 1e-4
    np.testing.assert_allclose(Q[-1], np.deg2rad(R.Q1_DEG), atol=1e-12)

"""Component variants inside SURVEY.md section 8(a) rows that round 1 left out: point gravity (row a4,
src/components.jl:89-98) and `LineForceWithMass` / `Spring(c, m)` (row a12, src/forces.jl:279-326)."""
import numpy as np
from scipy.integrate import solve_ivp

from multibody_jl_b200 import api


def test_point_gravity_reference_known_answer_and_kepler():
    s = api.multibody("point_gravity")
    assert s.n == 24
    f = lambda t, u: s.eval_rhs(u, s.p0, t)
    sol = solve_ivp(f, (0.0, 5.0), s.u0, method="DOP853", rtol=1e-10, atol=1e-12)
    r2 = np.array([sol.y[s.state_index(f"body2.r_0[{i+1}]"), -1] for i in range(3)])
    assert np.abs(r2 - [0.7867717, 0.478463, 0.0]).max() < 1e-1            # test/basic_tests.jl:168 (atol 1e-1)
    # independent: a point mass in -mu r / |r|^3 (the bodies do not interact; orientation plays no part)
    for k, (r0, v0) in enumerate((((0.0, 0.6, 0.0), (1.0, 0.0, 0.0)), ((0.6, 0.6, 0.0), (0.6, 0.0, 0.0)))):
        kep = solve_ivp(lambda t, y: np.concatenate([y[3:], -y[:3] / np.linalg.norm(y[:3]) ** 3]), (0.0, 5.0), np.array(r0 + v0),
                        method="DOP853", rtol=1e-11, atol=1e-13)
        got = np.array([sol.y[s.state_index(f"body{k+1}.r_0[{i+1}]"), -1] for i in range(3)])
        assert np.abs(got - kep.y[:3, -1]).max() < 1e-7
    assert np.abs(r2 - [0.7867717, 0.478463, 0.0]).max() < 2e-3            # (the printed digits are a default-tolerance Rodas4 run)


def test_spring_with_point_mass_closed_form():
    s = api.multibody("spring_with_mass")
    rng = np.random.default_rng(4)
    g = 9.80665
    for _ in range(6):
        c, ms, M = rng.uniform(10, 80), rng.uniform(0.0, 2.0), rng.uniform(0.5, 3.0)
        sj, vj = rng.uniform(0.05, 0.8), rng.uniform(-1, 1)
        du = s.eval_rhs(np.array([sj, vj]), np.array([c, ms, M]), 0.0)
        want = (M * g + ms * g / 2 - c * (sj - 0.1)) / (M + ms / 4)
        assert abs(du[0] - vj) < 1e-14 and abs(du[1] - want) < 1e-12 * max(1.0, abs(want))


def test_quaternion_and_euler_spherical_joints_agree():
    """test/basic_tests.jl:1170-1227 (`Spherical(state=true, quat=true)` vs `quat=false`, body.r_0 on 0:0.1:1, rtol 1e-5);
    also released with an out-of-plane spin so that all three rotational degrees of freedom move."""
    q, e = api.multibody("spherical_pendulum_quat"), api.multibody("spherical_pendulum_euler")
    assert q.n == 7 and e.n == 6
    tv = np.linspace(0.0, 1.0, 11)
    for spin in (0.0, 1.5):
        uq, ue = np.array(q.u0, float), np.array(e.u0, float)
        uq[q.state_index("joint.w_rel[2]")] = spin            # w_rel about the local y axis ...
        ue[e.state_index("joint.phid[2]")] = spin             # ... equals phid[2] at phi = 0
        out = []
        for s, u0 in ((q, uq), (e, ue)):
            sol = solve_ivp(lambda t, u: s.eval_rhs(u, s.p0, t), (0.0, 1.0), u0, method="DOP853", rtol=1e-10, atol=1e-12, t_eval=tv)
            names = s.observed_names
            obs = np.array([s.g.evaluate([s.observed[k] for k in names], {**{("u", i): sol.y[i, j] for i in range(s.n)}, ("t", 0): tv[j]}, s.tables)
                            for j in range(len(tv))])
            out.append(obs)
        assert np.abs(out[0] - out[1]).max() < 1e-7, spin
        assert abs(np.linalg.norm(out[0][-1]) - 1.0) < 1e-8     # the rod keeps its length (basic_tests.jl:621-622)

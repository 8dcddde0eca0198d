"""Generates the committed golden vectors under tests/golden/ (run once, in the build
container; scipy is the independent integrator):

    python tests/golden/make_golden.py

 * pendulum_dop853.npz      C1 angles, closed-form ODE of SURVEY.md B.0 (independent of the
                            code generator), scipy DOP853 rtol=atol=1e-13
 * spring_damper_dop853.npz C2 parameter sets, closed-form point-mass equations derived by
                            hand from docs/src/examples/spring_damper_system.md:21-61
 * robot6dof_radau.npz      C4 parameter sets, generated RHS/Jacobian, scipy Radau rtol=1e-10
 * reference_known_answers.json  numbers asserted by the reference's own tests (Appendix D)
The Julia reference cannot run here, so these pin the oracle's *integrators* against an
independent implementation and its *models* against hand-derived closed forms.
"""
import json
import os
import sys

import numpy as np
from scipy.integrate import solve_ivp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import mbpkg  # noqa: E402

mbpkg.load()
import mbo  # noqa: E402
from multibody_jl_b200 import ensembles  # noqa: E402
from multibody_jl_b200.mechanics.zoo import get_model  # noqa: E402

G = 9.80665


def pendulum():
    spec = ensembles.make_ensemble("C1", 1024)
    idx = np.array([0, 1, 100, 255, 256, 511, 777, 1023])
    m, ell, I33 = 1.0, 0.5, 0.001
    out = []
    for i in idx:
        f = lambda t, y: [y[1], -(m * G * ell * np.cos(y[0])) / (I33 + m * ell * ell)]
        r = solve_ivp(f, spec.tspan, spec.u0[:, i], method="DOP853", rtol=1e-13, atol=1e-13, t_eval=spec.saveat)
        out.append(r.y.T)
    np.savez(os.path.join(HERE, "pendulum_dop853.npz"), index=idx, saveat=spec.saveat, u=np.array(out))


def spring_damper():
    """body1: point mass m1 at r (r_cm = 0, so no torque: the Euler angles stay 0), pulled to
    bar1 = (0.3,0,0) by c1 (|r - bar1| - s0) + d1 d|r - bar1|/dt; body2: s'' = g - c2 (s - s0)/m2
    along -y.  Golden state order = the model's: r_0[3], v_0[3], phi[3], phid[3], p2.s, p2.v."""
    spec = ensembles.make_ensemble("C2", 6)
    sysm = get_model("spring_damper")
    P = {k: spec.p[sysm.param_index(k)] for k in sysm.param_names}
    out = []
    for i in range(6):
        c1, c2, d1, m1, m2 = (P[k][i] for k in ("spring1.c", "spring2.c", "damper1.d", "body1.m", "body2.m"))
        bar1 = np.array([0.3, 0.0, 0.0])

        def f(t, y):
            r, v, s, sd = y[0:3], y[3:6], y[6], y[7]
            rr = r - bar1
            L = max(np.linalg.norm(rr), 1e-10)
            e = rr / L
            fs = c1 * (L - 0.1) + d1 * np.dot(e, v)
            a = np.array([0.0, -G, 0.0]) - fs * e / m1
            return np.concatenate([v, a, [sd, G - c2 * (s - 0.1) / m2]])
        y0 = np.concatenate([spec.u0[0:6, i], spec.u0[12:14, i]])
        r = solve_ivp(f, spec.tspan, y0, method="DOP853", rtol=1e-13, atol=1e-13, t_eval=spec.saveat)
        full = np.zeros((len(spec.saveat), 14))
        full[:, 0:6] = r.y[0:6].T
        full[:, 12:14] = r.y[6:8].T
        out.append(full)
    np.savez(os.path.join(HERE, "spring_damper_dop853.npz"), index=np.arange(6), saveat=spec.saveat, u=np.array(out))


def robot():
    idx = np.array([0, 1, 2, 700001])
    sysm = get_model("robot6dof")
    tb = sysm.tables.pack()
    out = []
    for i in idx:
        spec = ensembles.make_ensemble("C4", 1, offset=int(i))
        u0, p = spec.u0[:, 0].copy(), spec.p[:, 0].copy()
        r = solve_ivp(lambda t, u: mbo.eval_fn("robot6dof", 0, u, p, t, tb), spec.tspan, u0, method="Radau",
                      jac=lambda t, u: mbo.jac_dense("robot6dof", u, p, t, tb), rtol=1e-10, atol=1e-12,
                      t_eval=spec.saveat)
        assert r.status == 0
        out.append(r.y.T)
        print("robot", i, r.nfev)
    np.savez(os.path.join(HERE, "robot6dof_radau.npz"), index=idx, saveat=spec.saveat, u=np.array(out))


def known_answers():
    ka = {
        "pendulum_phi_range": {"min": -np.pi, "max": 0.0, "rtol": 1e-2, "atol": 1e-2, "src": "test/basic_tests.jl:104-105"},
        "pendulum_damper_power_t1_d10": {"value": -1.94758, "atol": 1e-2, "src": "test/basic_tests.jl:118"},
        "damped_pendulum_end": {"value": -np.pi / 2, "rtol": 1e-2, "src": "test/basic_tests.jl:201"},
        "robot_axis1_tracking_error": {"bound": 0.01, "grid": "0:0.1:2", "src": "test/test_robot.jl:320-325"},
        "robot_q1_deg": [60.0, -70.0, -35.0, 45.0, 110.0, 45.0],
        "spring_damper_velocity_settles": {"src": "test/basic_tests.jl:420"},
    }
    json.dump(ka, open(os.path.join(HERE, "reference_known_answers.json"), "w"), indent=1)


if __name__ == "__main__":
    pendulum()
    spring_damper()
    robot()
    known_answers()

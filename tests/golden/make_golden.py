"""Generates the committed golden vectors under tests/golden/ (run once, in the build
container; scipy is the independent integrator):

    python tests/golden/make_golden.py

 * pendulum_dop853.npz      C1 angles, closed-form ODE of SURVEY.md B.0 (independent of the
                            code generator), scipy DOP853 rtol=atol=1e-13
 * spring_damper_dop853.npz C2 parameter sets, closed-form point-mass equations derived by
                            hand from docs/src/examples/spring_damper_system.md:21-61
 * robot6dof_radau.npz      C4 parameter sets, generated RHS/Jacobian, scipy Radau rtol=1e-10
 * fourbar_lagrange.npz      C3 crank speeds, independent 1-DOF Lagrange derivation (numpy forward
                            kinematics + Newton closure), scipy DOP853 rtol=1e-7
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


def _rot(axis, a):
    c, s_ = np.cos(a), np.sin(a)
    if axis == 0:
        return np.array([[1, 0, 0], [0, c, -s_], [0, s_, c]])
    if axis == 1:
        return np.array([[c, 0, s_], [0, 1, 0], [-s_, 0, c]])
    return np.array([[c, -s_, 0], [s_, c, 0], [0, 0, 1]])


class FourbarLagrange:
    """Independent 1-DOF derivation of the four-bar of test/test_fourbar.jl:13-42: closure
    solved by Newton for q_D(phi), then  m_red(phi) phi'' + m_red'(phi) phi'^2 / 2 + V'(phi) = 0
    with m_red, V from plain forward kinematics and nested central differences."""

    def __init__(self):
        self.qd = np.array([-0.2, 0.0, 0.0, 0.0, 0.0])      # s, rev, rev1, j3, j4

    @staticmethod
    def fk(phi, qd):
        s, a_rev, a_rev1, a3, a4 = qd
        R1 = _rot(0, phi)
        p2 = np.array([1.2 + s, 0.0, 0.0])
        p3 = p2 + np.array([0.0, 0.2, 0.0])
        T3 = _rot(1, a_rev) @ _rot(2, a_rev1)
        TA = R1 @ _rot(0, a3) @ _rot(1, a4)
        pA = R1 @ np.array([0.0, 0.5, 0.1])
        pB = p3 + T3 @ np.array([-1.0, 0.3, 0.1])
        nb = TA.T @ T3 @ np.array([0.0, 0.0, 1.0])
        g = np.concatenate([pB - pA, nb[:2]])
        coms = [R1 @ np.array([0.0, 0.25, 0.05]), p2 + np.array([0.0, 0.1, 0.0]), p3 + T3 @ np.array([-0.5, 0.15, 0.05])]
        return g, coms, [R1, np.eye(3), T3]

    def solve_qd(self, phi, guess=None):
        qd = np.array(self.qd if guess is None else guess, dtype=float)
        for _ in range(60):
            g = self.fk(phi, qd)[0]
            if np.abs(g).max() < 1e-14:
                break
            J = np.zeros((5, 5))
            for j in range(5):
                e = np.zeros(5); e[j] = 1e-7
                J[:, j] = (self.fk(phi, qd + e)[0] - self.fk(phi, qd - e)[0]) / 2e-7
            qd = qd - np.linalg.solve(J, g)
        return qd

    def energy_terms(self, phi):
        """(m_red, V) at phi."""
        h = 1e-5
        q0 = self.solve_qd(phi)
        self.qd = q0
        qp, qm = self.solve_qd(phi + h, q0), self.solve_qd(phi - h, q0)
        _, c0, T0 = self.fk(phi, q0)
        _, cp, Tp = self.fk(phi + h, qp)
        _, cm, Tm = self.fk(phi - h, qm)
        mred = 0.0
        for k in range(3):
            dr = (cp[k] - cm[k]) / (2 * h)
            Wm = ((Tp[k] - Tm[k]) / (2 * h)) @ T0[k].T
            w = np.array([Wm[2, 1], Wm[0, 2], Wm[1, 0]])
            mred += 1.0 * dr @ dr + 0.001 * w @ w
        V = sum(1.0 * G * c[1] for c in c0)
        return mred, V

    def rhs(self, t, y):
        phi, w = y
        h = 1e-4
        m0, _ = self.energy_terms(phi)
        mp, Vp = self.energy_terms(phi + h)
        mm, Vm = self.energy_terms(phi - h)
        self.energy_terms(phi)          # leave the Newton guess at phi
        return [w, -(0.5 * (mp - mm) / (2 * h) * w * w + (Vp - Vm) / (2 * h)) / m0]


def fourbar():
    w0s = np.array([5.235987755982989, 3.0, 6.5])
    sv = np.linspace(0.0, 1.4399, 21)[1:]
    out = []
    for w0 in w0s:
        fb = FourbarLagrange()
        r = solve_ivp(fb.rhs, (0.0, 1.4399), [0.0, w0], method="DOP853", rtol=1e-7, atol=1e-9, t_eval=sv)   # the nested differences carry ~1e-7 noise
        fb2 = FourbarLagrange()
        rows = []
        for k in range(len(sv)):
            qd = fb2.solve_qd(r.y[0, k])
            fb2.qd = qd
            rows.append(np.concatenate([r.y[:, k], qd]))
        out.append(np.array(rows))
        print("fourbar", w0, r.nfev, rows[-1])
    np.savez(os.path.join(HERE, "fourbar_lagrange.npz"), w0=w0s, saveat=sv, u=np.array(out))


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
    if len(sys.argv) > 1:
        for name in sys.argv[1:]:
            globals()[name]()
        sys.exit(0)
    fourbar()
    pendulum()
    spring_damper()
    robot()
    known_answers()

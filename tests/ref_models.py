"""Second, independent derivation of the Robot6DOF right-hand side: a
plain-numpy recursive Newton-Euler pass in WORLD coordinates (the product's code generator
works in local frames on a symbolic DAG, mechanics/mechanism.py).  Data restated from
/root/reference/src/robot/robot_components.jl:432-570 and FullRobot.jl:121-246."""
import numpy as np

G = np.array([0.0, -9.80665, 0.0])
AXES_N = [(0, 1, 0), (1, 0, 0), (1, 0, 0), (0, 1, 0), (1, 0, 0), (0, 1, 0)]
# r (a->b), r_cm, m, (I11, I22, I33, I21)
LINKS = [
    ((0, 0.324, 0.3), None, 1.0, (0.001, 1.16, 0.001, 0.0)),
    ((0, 0.65, 0), (0.172, 0.205, 0), 56.5, (2.58, 0.64, 2.73, -0.46)),
    ((0, 0.414, -0.155), (0.064, -0.034, 0), 26.4, (0.279, 0.245, 0.413, -0.070)),
    ((0, 0.186, 0), None, 28.7, (1.67, 0.081, 1.67, 0.0)),
    ((0, 0.125, 0), None, 5.2, (1.25, 0.81, 1.53, 0.0)),
    ((0, 0, 0), (0.05, 0.05, 0.05), 0.5, (0.001, 0.001, 0.001, 0.0)),
]
BASE_R = np.array([0.0, 0.351, 0.0])
R_LOAD = np.array([0.1, 0.25, 0.1])
# type, ratio, w, D, J, k, c, cd
DRIVES = [(1, -105.0, 4590.0, 0.6, 0.0013, 1.1616, 43.0, 0.005), (1, 210.0, 5500.0, 0.6, 0.0013, 1.1616, 8.0, 0.01),
          (1, 60.0, 5500.0, 0.6, 0.0013, 1.1616, 58.0, 0.04), (2, -99.0, 6250.0, 0.55, 1.6e-4, 0.2365, 0, 0),
          (2, 79.2, 6250.0, 0.55, 1.8e-4, 0.2608, 0, 0), (2, -99.0, 7400.0, 0.27, 4.3e-5, 0.0842, 0, 0)]


def rodrigues(n, phi):
    n = np.asarray(n, float)
    K = np.array([[0, -n[2], n[1]], [n[2], 0, -n[0]], [-n[1], n[0], 0]])
    return np.eye(3) + np.sin(phi) * K + (1 - np.cos(phi)) * K @ K


def inverse_dynamics(q, qd, qdd, m_load):
    """Joint torques of the open chain for given motion (world-frame Newton-Euler)."""
    T = np.eye(3)
    p = BASE_R.copy()
    om = np.zeros(3); al = np.zeros(3); ap = np.zeros(3)
    bodies = []          # (joint index, F, N, com)
    jp = []; ja = []
    for i in range(6):
        a = T @ np.array(AXES_N[i], float)
        al = al + a * qdd[i] + np.cross(om, a * qd[i])
        om = om + a * qd[i]
        T = T @ rodrigues(AXES_N[i], q[i])
        jp.append(p.copy()); ja.append(a)
        r, rcm, m, (I11, I22, I33, I21) = LINKS[i]
        r = np.array(r, float)
        rcm = r / 2 if rcm is None else np.array(rcm, float)
        items = [(m, rcm, np.array([[I11, I21, 0], [I21, I22, 0], [0, 0, I33]]))]
        if i == 5:
            items.append((m_load, R_LOAD, np.eye(3) * 0.001))
        for (mb, rc, I) in items:
            d = T @ rc
            acom = ap + np.cross(al, d) + np.cross(om, np.cross(om, d))
            Iw = T @ I @ T.T
            bodies.append((i, mb * (acom - G), Iw @ al + np.cross(om, Iw @ om), p + d))
        d = T @ r
        ap = ap + np.cross(al, d) + np.cross(om, np.cross(om, d))
        p = p + d
    tau = np.zeros(6)
    for i in range(6):
        for (j, F, N, c) in bodies:
            if j >= i:
                tau[i] += ja[i] @ (N + np.cross(c - jp[i], F))
    return tau


def robot_rhs(u, p, q_ref, qd_ref):
    """du/dt in the product's state order (mechanics/robot.py docstring)."""
    q = u[0:12:2]; w = u[1:12:2]
    m_load = p[0]
    kp, ks, Ts = p[1::3], p[2::3], p[3::3]
    du = np.zeros(36)
    h = inverse_dynamics(q, w, np.zeros(6), m_load)
    M = np.column_stack([inverse_dynamics(q, w, e, m_load) - h for e in np.eye(6)])
    tau = np.zeros(6)
    k0 = 12
    for i, (typ, ratio, wm_, D, J, k, c, cd) in enumerate(DRIVES):
        La = 250.0 / (2 * D * wm_); Cc = 0.004 * D / wm_
        if typ == 1:
            phim, wmot, cur, vC, x = u[k0:k0 + 5]
        else:
            cur, vC, x = u[k0:k0 + 3]
            phim, wmot = ratio * q[i], ratio * w[i]
        e = ratio * qd_ref[i] + kp[i] * (ratio * q_ref[i] - phim) - wmot
        iref = ks[i] * (x + e)
        if typ == 1:
            tsd = c * (ratio * q[i] - phim) + cd * (ratio * w[i] - wmot)
            du[k0] = wmot; du[k0 + 1] = (k * cur + tsd) / J
            du[k0 + 2] = (-5 * vC - 250 * cur - k * wmot) / La; du[k0 + 3] = (cur - iref) / (10 * Cc); du[k0 + 4] = e / Ts[i]
            tau[i] = -ratio * tsd
            k0 += 5
        else:
            du[k0] = (-5 * vC - 250 * cur - k * wmot) / La; du[k0 + 1] = (cur - iref) / (10 * Cc); du[k0 + 2] = e / Ts[i]
            tau[i] = ratio * k * cur
            M[i, i] += J * ratio * ratio
            k0 += 3
    qdd = np.linalg.solve(M, tau - h)
    du[0:12:2] = w
    du[1:12:2] = qdd
    return du


# ---- rope (C5): independent forward kinematics for the energy check -------------------------
def _axis_rot_T(k, a):
    """local->world elementary rotation about axis k (1-based)."""
    return rodrigues([1.0 if i == k - 1 else 0.0 for i in range(3)], a)


def rope_energy(u, m_body, n_links=6, l=1.0, m_rope=6.0, dirv=(0.05, 1.0, 0.0)):
    """Total mechanical energy of the stiff rope of docs/src/examples/ropes_and_cables.md:24-36
    from the Euler-angle states (sequence [1,2,3]); velocities by differencing the forward
    kinematics along the state velocity, so nothing is shared with the product's equations."""
    dirv = np.asarray(dirv, float) / np.linalg.norm(dirv)
    li, mi = l / n_links, m_rope / n_links

    def fk(phis):
        T = np.eye(3)
        p = np.zeros(3)
        out = []
        for i in range(n_links + 1):
            a = phis[3 * i:3 * i + 3]
            # R = A3 A2 A1 is world->local (orientation.jl:288-299), so local->world appends A1' A2' A3'
            T = T @ _axis_rot_T(1, a[0]) @ _axis_rot_T(2, a[1]) @ _axis_rot_T(3, a[2])
            if i < n_links:
                out.append((mi, p + T @ (li / 2 * dirv), T.copy()))
                p = p + T @ (li * dirv)
            else:
                out.append((m_body, p.copy(), T.copy()))
        return out
    phi = np.concatenate([u[6 * i:6 * i + 3] for i in range(n_links + 1)])
    phid = np.concatenate([u[6 * i + 3:6 * i + 6] for i in range(n_links + 1)])
    eps = 1e-6
    b0, bp, bm = fk(phi), fk(phi + eps * phid), fk(phi - eps * phid)
    E = 0.0
    for (m, c, T), (_, cp, Tp), (_, cm, Tm) in zip(b0, bp, bm):
        v = (cp - cm) / (2 * eps)
        Wm = ((Tp - Tm) / (2 * eps)) @ T.T
        w = np.array([Wm[2, 1], Wm[0, 2], Wm[1, 0]])
        E += 0.5 * m * v @ v + 0.5 * 0.001 * w @ w + m * 9.80665 * c[1]
    return E


def elastic_rope_energy(u, names, p, pnames, n_links=3, l=1.0, m_rope=5.0, dirv=(0.0, -1.0, 0.0)):
    """Total mechanical energy (kinetic + gravity + link springs) of the flexible rope of
    test/basic_tests.jl:928-959 from its named states; velocities by differencing an independent
    numpy forward kinematics along the state velocity."""
    dirv = np.asarray(dirv, float) / np.linalg.norm(dirv)
    li, mi = l / n_links, m_rope / n_links
    ix = {nm: k for k, nm in enumerate(names)}
    par = {nm: p[k] for k, nm in enumerate(pnames)}
    ci = n_links * par["rope.c"]

    def unpack(vec_of):
        phis = [[vec_of(f"rope.joint_{i+1}.phi[{k+1}]", f"rope.joint_{i+1}.phid[{k+1}]") for k in range(3)] for i in range(n_links + 1)]
        ss = [vec_of(f"rope.flexibility_{i+1}.s", f"rope.flexibility_{i+1}.v") for i in range(n_links)]
        return phis, ss

    def fk(eps):
        phis, ss = unpack(lambda q, v: u[ix[q]] + eps * u[ix[v]])
        T, pos, out = np.eye(3), np.zeros(3), []
        for i in range(n_links + 1):
            a = phis[i]
            T = T @ _axis_rot_T(1, a[0]) @ _axis_rot_T(2, a[1]) @ _axis_rot_T(3, a[2])
            if i < n_links:
                out.append((mi, pos + T @ (li / 2 * dirv), T.copy()))
                pos = pos + T @ (ss[i] * dirv)
            else:
                out.append((par["body.m"], pos.copy(), T.copy()))
        return out, ss
    eps = 1e-6
    (b0, s0), (bp, _), (bm, _) = fk(0.0), fk(eps), fk(-eps)
    E = sum(0.5 * ci * (s - li) ** 2 for s in s0)
    for (m, c, T), (_, cp, Tp), (_, cm, Tm) in zip(b0, bp, bm):
        v = (cp - cm) / (2 * eps)
        Wm = ((Tp - Tm) / (2 * eps)) @ T.T
        w = np.array([Wm[2, 1], Wm[0, 2], Wm[1, 0]])
        E += 0.5 * m * v @ v + 0.5 * 0.001 * w @ w + m * 9.80665 * c[1]
    return E

"""Independent derivation of the robot's drive train (the reduced motor / controller equations of
mechanics/robot.py were only ever compared with a copy of themselves).

Here the reference's `Motor` is taken as what it is -- a NETLIST (src/robot/robot_components.jl:336-408: eight
resistors, C, La, three ideal op-amps, an EMF, two controlled voltage sources, five grounds; every `connect` below is
transcribed from :373-411) -- and solved by modified nodal analysis for d(La.i)/dt and d(C.v)/dt; `Controller`
(:190-219) is evaluated as the block diagram it is (Gain, Feedback, Add3(k3 = -1), PI).  Nothing here looks at
mechanics/robot.py.  The resulting derivatives must equal the right-hand-side rows of the robot6dof model at random
states, for all six axes (data: src/robot/FullRobot.jl:186-246)."""
import numpy as np

from multibody_jl_b200.mechanics.zoo import get_model

# (type, ratio, w, D, J, k) per axis, FullRobot.jl:186-246
AXES = [(1, -105.0, 4590.0, 0.6, 0.0013, 1.1616), (1, 210.0, 5500.0, 0.6, 0.0013, 1.1616), (1, 60.0, 5500.0, 0.6, 0.0013, 1.1616),
        (2, -99.0, 6250.0, 0.55, 1.6e-4, 0.2365), (2, 79.2, 6250.0, 0.55, 1.8e-4, 0.2608), (2, -99.0, 7400.0, 0.27, 4.3e-5, 0.0842)]

CONNECT = [("La.n", "emf.p"), ("Ra.n", "La.p"), ("Rd2.n", "diff.n1"), ("C.n", "OpI.p2"), ("OpI.p2", "power.p1"), ("Vs.p", "Rd2.p"),
           ("diff.n1", "Rd1.p"), ("Rd1.n", "diff.p2"), ("diff.p2", "Ri.p"), ("Ri.n", "OpI.n1"), ("OpI.n1", "C.p"), ("power.n1", "Rp1.p"),
           ("power.p2", "Rp1.n"), ("Rp1.p", "Rp2.p"), ("power.p2", "Ra.p"), ("Rd3.p", "hall2.p"), ("Rd3.n", "diff.p1"), ("Rd3.n", "Rd4.p"),
           ("Vs.n", "g1.g"), ("g2.g", "hall2.n"), ("Rd4.n", "g3.g"), ("g3.g", "OpI.p1"), ("g5.g", "Rp2.n"), ("emf.n", "hall1.p"),
           ("hall1.n", "g4.g"), ("OpI.n2", "power.n2"), ("OpI.p1", "OpI.n2"), ("OpI.p1", "diff.n2")]
RES = {"Ra": 250.0, "Rd2": 100.0, "Ri": 10.0, "Rd1": 100.0, "Rp1": 200.0, "Rp2": 50.0, "Rd4": 100.0, "Rd3": 100.0}
TWO = list(RES) + ["C", "La", "Vs", "hall2", "hall1", "emf"]           # one-ports: pins p, n; unknown current i = p.i
OPAMPS = ["power", "diff", "OpI"]                                       # pins p1, n1 (inputs), p2, n2 (output port)
GROUNDS = ["g1", "g2", "g3", "g4", "g5"]


def motor_derivatives(k, wD, iL, vC, i_ref, w_rotor):
    """(d La.i/dt, d C.v/dt, current through hall1) of the netlist for the given states / inputs."""
    w, D = wD
    L, Cc = 250.0 / (2.0 * D * w), 0.004 * D / w                       # robot_components.jl:343,346
    pins = [f"{e}.{q}" for e in TWO for q in "pn"] + [f"{o}.{q}" for o in OPAMPS for q in ("p1", "n1", "p2", "n2")] + [f"{g}.g" for g in GROUNDS]
    parent = {p: p for p in pins}

    def find(a):
        while parent[a] != a:
            a = parent[a]
        return a
    for a, b in CONNECT:
        parent[find(a)] = find(b)
    nodes = sorted({find(p) for p in pins})
    nid = {r: i for i, r in enumerate(nodes)}
    node = lambda p: nid[find(p)]
    nn, ne = len(nodes), len(TWO)
    cur = {e: nn + i for i, e in enumerate(TWO)}
    out = {o: nn + ne + i for i, o in enumerate(OPAMPS)}
    nu = nn + ne + len(OPAMPS)
    rows, rhs = [], []

    def eq(coeffs, b=0.0):
        r = np.zeros(nu)
        for j, c in coeffs:
            r[j] += c
        rows.append(r); rhs.append(b)
    ground_nodes = {node(f"{g}.g") for g in GROUNDS}
    for gnode in ground_nodes:
        eq([(gnode, 1.0)])
    # Kirchhoff's current law at every other node: sum of the currents flowing INTO the connected components = 0
    inc = {i: [] for i in range(nn)}
    for e in TWO:
        inc[node(f"{e}.p")].append((cur[e], 1.0)); inc[node(f"{e}.n")].append((cur[e], -1.0))
    for o in OPAMPS:                                                     # inputs draw nothing; output port: p2.i + n2.i = 0
        inc[node(f"{o}.p2")].append((out[o], 1.0)); inc[node(f"{o}.n2")].append((out[o], -1.0))
    for i in range(nn):
        if i not in ground_nodes:
            eq(inc[i])
    v = lambda e: [(node(f"{e}.p"), 1.0), (node(f"{e}.n"), -1.0)]
    for e, R in RES.items():
        eq(v(e) + [(cur[e], -R)])                                        # v = R i
    eq(v("C"), vC)                                                        # capacitor: its voltage is the state
    eq([(cur["La"], 1.0)], iL)                                            # inductor: its current is the state
    eq(v("Vs"), i_ref)                                                    # Vs.v = current_ref (convert2, gain 1)
    eq(v("hall2") + [(cur["hall1"], -1.0)])                               # hall2.v = hall1.i (convert1, gain 1)
    eq(v("hall1"))                                                        # current sensor: no voltage drop
    eq(v("emf"), k * w_rotor)                                             # EMF: v = k w
    for o in OPAMPS:
        eq([(node(f"{o}.p1"), 1.0), (node(f"{o}.n1"), -1.0)])            # ideal op-amp: v1 = 0
    A, b = np.array(rows), np.array(rhs)
    assert A.shape == (nu, nu), A.shape
    x = np.linalg.solve(A, b)
    vLa = x[node("La.p")] - x[node("La.n")]
    return vLa / L, x[cur["C"]] / Cc, x[cur["hall1"]]


def controller(ratio, kp, ks, Ts, q_ref, qd_ref, motor_angle, motor_speed, x_pi):
    """Controller (:190-219) as its block diagram: returns (current_ref, d PI.x/dt)."""
    gain1 = ratio * q_ref                          # gain1.input = angle_ref
    gain2 = ratio * qd_ref                         # gain2.input = speed_ref
    feedback1 = gain1 - motor_angle                # Feedback: input1 - input2
    P = kp * feedback1
    e = 1.0 * gain2 + 1.0 * P + (-1.0) * motor_speed   # Add3(k3 = -1): input1 = gain2, input2 = P, input3 = motorSpeed
    return ks * (x_pi + e), e / Ts                 # Blocks.PI(k, T): D(x) = err / T, output = k (x + err)


def test_motor_netlist_and_controller_blocks_equal_the_model_rows():
    s = get_model("robot6dof")
    rng = np.random.default_rng(11)
    ix = s.state_index
    for trial in range(5):
        u = np.array(s.u0, float) + 0.3 * rng.standard_normal(s.n)
        p = np.array(s.p0, float) * (1.0 + 0.3 * rng.random(s.m))
        t = float(rng.uniform(0.05, 1.9))
        du = s.eval_rhs(u, p, t)
        for a, (typ, ratio, w, D, J, k) in enumerate(AXES):
            ax = f"axis{a+1}"
            kp, ks, Ts = (p[s.param_index(f"{ax}.controller.{q}")] for q in ("kp", "ks", "Ts"))
            q_ref, qd_ref = s.tables.eval("cubic", a, 0, t), s.tables.eval("linear", 6 + a, 0, t)
            if typ == 1:
                ang, spd = u[ix(f"{ax}.motor.Jmotor.phi")], u[ix(f"{ax}.motor.Jmotor.w")]
            else:                                  # AxisType2: rotor rigidly geared to the joint (IdealGear, :221-262): phi_a = ratio phi_b
                ang, spd = ratio * u[ix(f"mechanics.r{a+1}.phi")], ratio * u[ix(f"mechanics.r{a+1}.w")]
            iL, vC, xpi = u[ix(f"{ax}.motor.La.i")], u[ix(f"{ax}.motor.C.v")], u[ix(f"{ax}.controller.PI.x")]
            i_ref, dx = controller(ratio, kp, ks, Ts, q_ref, qd_ref, ang, spd, xpi)
            diL, dvC, ihall = motor_derivatives(k, (w, D), iL, vC, i_ref, spd)
            assert abs(ihall - iL) < 1e-9 * max(1.0, abs(iL))              # the measured current is the armature current
            for name, want in ((f"{ax}.motor.La.i", diL), (f"{ax}.motor.C.v", dvC), (f"{ax}.controller.PI.x", dx)):
                got = du[ix(name)]
                assert abs(got - want) <= 1e-9 * max(1.0, abs(want)), (trial, name, got, want)


def test_rotor_sees_the_emf_torque_k_i():
    """EMF (:341): flange.tau = -k i flows into the EMF, so +k i acts on Jmotor; for AxisType1 d(Jmotor.w)/d(La.i) = k / J."""
    s = get_model("robot6dof")
    u, p = np.array(s.u0, float), np.array(s.p0, float)
    for a, (typ, ratio, w, D, J, k) in enumerate(AXES[:3]):
        up = u.copy()
        up[s.state_index(f"axis{a+1}.motor.La.i")] += 1.0
        row = s.state_index(f"axis{a+1}.motor.Jmotor.w")
        assert abs((s.eval_rhs(up, p, 0.4)[row] - s.eval_rhs(u, p, 0.4)[row]) - k / J) < 1e-8 * k / J

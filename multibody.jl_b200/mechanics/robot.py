"""Robot6DOF: the north-star model (reference src/robot/FullRobot.jl:78-266).

Mechanics: `MechanicalStructure` (src/robot/robot_components.jl:415-573).
Axes: `AxisType1` (:143-188, axes 1-3, compliant gear) and `AxisType2`
(:98-140, axes 4-6, rigid ideal gear), `Motor` (:324-411), `Controller`
(:190-219), reference path `KinematicPTP` (src/robot/path_planning.jl:177-214)
over `point_to_point` (src/robot/ptp.jl:37-143).
"""
from __future__ import annotations

import math
from typing import Sequence

import numpy as np

from .builder import ModelBuilder, ReducedSystem
from .mechanism import Mechanism


def point_to_point(time: Sequence[float], q0, q1, qd_max, qdd_max, t0: float = 0.0):
    """Synchronised trapezoidal PTP profile (restates src/robot/ptp.jl:37-143).
    Returns (Q, Qd, Qdd, t_end) with Q of shape (len(time), naxis)."""
    q0 = np.atleast_1d(np.asarray(q0, dtype=float))
    q1 = np.atleast_1d(np.asarray(q1, dtype=float))
    qd_max = np.broadcast_to(np.asarray(qd_max, dtype=float), q0.shape)
    qdd_max = np.broadcast_to(np.asarray(qdd_max, dtype=float), q0.shape)
    dq = q1 - q0
    sd_max_inv = np.max(np.abs(dq / qd_max))
    sdd_max_inv = np.max(np.abs(dq / qdd_max))
    time = np.asarray(time, dtype=float)
    nt = len(time)
    if sd_max_inv <= np.finfo(float).eps or sdd_max_inv <= np.finfo(float).eps:
        z = np.zeros((nt, len(q0)))
        return z + q0, z.copy(), z.copy(), float(time[0])
    sd_max = 1.0 / sd_max_inv
    sdd_max = 1.0 / sdd_max_inv
    Ta1 = math.sqrt(1.0 / sdd_max)
    Ta2 = sd_max / sdd_max
    noW = Ta2 >= Ta1
    Tv = Ta1 if noW else 1.0 / sd_max
    Te = 2.0 * Ta1 if noW else Tv + Ta2
    Ta1s, Ta2s, Tvs, Tes = Ta1 + t0, Ta2 + t0, Tv + t0, Te + t0
    sd_max2 = sdd_max * Ta1
    s1 = sdd_max * (Ta1 * Ta1 if noW else Ta2 * Ta2) / 2.0
    if noW:
        s2 = s1 + sd_max2 * (Te - Ta1) - (sdd_max / 2.0) * (Te - Ta1) ** 2
    else:
        s2 = s1 + sd_max * (Tv - Ta2)
    s3 = s2 + sd_max * (Te - Tv) - (sdd_max / 2.0) * (Te - Tv) * (Te - Tv)
    Q = np.zeros((nt, len(q0))); Qd = np.zeros_like(Q); Qdd = np.zeros_like(Q)
    for i, t in enumerate(time):
        sd = sdd = 0.0
        if t < t0:
            s = 0.0
        elif noW:
            if t < Ta1s:
                s = (sdd_max / 2.0) * (t - t0) ** 2; sd = sdd_max * (t - t0); sdd = sdd_max
            elif t < Tes:
                s = s1 + sd_max2 * (t - Ta1s) - (sdd_max / 2.0) * (t - Ta1s) ** 2
                sd = sd_max2 + (Ta1s - t) * sdd_max; sdd = -sdd_max
            else:
                s = s2
        elif t < Ta2s:
            s = (sdd_max / 2.0) * (t - t0) ** 2; sd = sdd_max * (t - t0); sdd = sdd_max
        elif t < Tvs:
            s = s1 + sd_max * (t - Ta2s); sd = sd_max
        elif t < Tes:
            s = s2 + sd_max * (t - Tvs) - (sdd_max / 2.0) * (t - Tvs) ** 2
            sd = sd_max + (Tvs - t) * sdd_max; sdd = -sdd_max
        else:
            s = s3
        Q[i] = q0 + dq * s; Qd[i] = dq * sd; Qdd[i] = dq * sdd
    return Q, Qd, Qdd, Tes


# FullRobot.jl:186-246 (axis data); AxisType defaults robot_components.jl:98-99,143-144
AXES = [
    # type, ratio,  w,    D,    J,      k,      c,   cd
    (1, -105.0, 4590.0, 0.6, 0.0013, 1.1616, 43.0, 0.005),
    (1, 210.0, 5500.0, 0.6, 0.0013, 1.1616, 8.0, 0.01),
    (1, 60.0, 5500.0, 0.6, 0.0013, 1.1616, 58.0, 0.04),
    (2, -99.0, 6250.0, 0.55, 1.6e-4, 0.2365, 0.0, 0.0),
    (2, 79.2, 6250.0, 0.55, 1.8e-4, 0.2608, 0.0, 0.0),
    (2, -99.0, 7400.0, 0.27, 4.3e-5, 0.0842, 0.0, 0.0),
]
Q0_DEG = [-60.0, 20.0, 90.0, 0.0, -110.0, 0.0]      # FullRobot.jl:147-152
Q1_DEG = [60.0, -70.0, -35.0, 45.0, 110.0, 45.0]    # FullRobot.jl:153-158
QD_MAX = [3.0, 1.5, 5.0, 3.1, 3.1, 4.1]             # FullRobot.jl:126
QDD_MAX = [15.0, 15.0, 15.0, 60.0, 60.0, 60.0]      # FullRobot.jl:127
JOINT_AXES = [(0, 1, 0), (1, 0, 0), (1, 0, 0), (0, 1, 0), (1, 0, 0), (0, 1, 0)]  # robot_components.jl:440-445

_I = lambda a=0.001, b=0.001, c=0.001, d=0.0: (a, b, c, d, 0.0, 0.0)
# (r, r_cm or None, m, I)  robot_components.jl:446-535
LINKS = [
    ((0, 0.351, 0), None, 1.0, _I()),                                        # b0
    ((0, 0.324, 0.3), None, 1.0, _I(b=1.16)),                                # b1
    ((0, 0.65, 0), (0.172, 0.205, 0), 56.5, _I(2.58, 0.64, 2.73, -0.46)),    # b2
    ((0, 0.414, -0.155), (0.064, -0.034, 0), 26.4, _I(0.279, 0.245, 0.413, -0.070)),  # b3
    ((0, 0.186, 0), None, 28.7, _I(1.67, 0.081, 1.67)),                      # b4
    ((0, 0.125, 0), None, 5.2, _I(1.25, 0.81, 1.53)),                        # b5
    ((0, 0, 0), (0.05, 0.05, 0.05), 0.5, _I()),                              # b6
]
R_LOAD = (0.1, 0.25, 0.1)  # FullRobot.jl:122


def robot6dof() -> ReducedSystem:
    """36-state reduced Robot6DOF (docs/src/examples/robot.md:135).

    State order (ours; MTK's is not observable here): per axis i=1..6
    [r_i.phi, r_i.w], then per axis the drive states: axes 1-3
    [Jmotor.phi, Jmotor.w, La.i, C.v, PI.x], axes 4-6 [La.i, C.v, PI.x].
    Tunable parameters: load.m, then kp, ks, Ts per axis."""
    mb = ModelBuilder("robot6dof")
    mech = Mechanism(mb)  # World(): g = 9.80665 along -y; the 9.81 argument is ignored (robot_components.jl:415-420)
    time = np.arange(401) * 0.01                      # path_planning.jl:29
    Q, Qd, _, _ = point_to_point(time, np.deg2rad(Q0_DEG), np.deg2rad(Q1_DEG), QD_MAX, QDD_MAX)
    mb.tables.set_grid(0.0, 0.01, 401)
    qtab = [mb.tables.add(Q[:, i], cubic=True) for i in range(6)]
    qdtab = [mb.tables.add(Qd[:, i], cubic=False) for i in range(6)]

    mLoad = mb.param("mechanics.load.m", 15.0)
    kp = []; ks = []; Ts = []
    for i in range(6):
        kp.append(mb.param(f"axis{i+1}.controller.kp", 5.0))     # FullRobot.jl:128-145
        ks.append(mb.param(f"axis{i+1}.controller.ks", 0.5))
        Ts.append(mb.param(f"axis{i+1}.controller.Ts", 0.05))

    q0 = np.deg2rad(Q0_DEG)
    phi = []; w = []
    for i in range(6):
        phi.append(mb.state(f"mechanics.r{i+1}.phi", q0[i]))    # test/test_robot.jl:273-278
        w.append(mb.state(f"mechanics.r{i+1}.w", 0.0))

    # mechanical chain (robot_components.jl:547-570)
    fr = mech.body_shape(mech.world(), LINKS[0][0], LINKS[0][2], LINKS[0][1], LINKS[0][3])
    joints = []
    for i in range(6):
        jb = mech.revolute(fr, JOINT_AXES[i], phi[i], w[i], name=f"r{i+1}")
        joints.append(jb)
        r, rcm, m, I = LINKS[i + 1]
        fr = mech.body_shape(jb, r, m, rcm, I)
    mech.body(fr, mLoad, R_LOAD, _I())                              # load, :536-545

    # drives
    for i, (typ, ratio, wm, D, J, k, c, cd) in enumerate(AXES):
        La = 250.0 / (2.0 * D * wm)                                 # Motor, :343
        C = 0.004 * D / wm                                          # :346
        q_ref = mb.cubic(qtab[i])
        qd_ref = mb.linear(qdtab[i])
        if typ == 1:
            phim = mb.state(f"axis{i+1}.motor.Jmotor.phi", q0[i] * ratio)   # test_robot.jl:280-286
            wmot = mb.state(f"axis{i+1}.motor.Jmotor.w", 0.0)
        else:
            phim = ratio * phi[i]                                    # IdealGear: phi_a = ratio phi_b
            wmot = ratio * w[i]
        cur = mb.state(f"axis{i+1}.motor.La.i", 0.0)
        vC = mb.state(f"axis{i+1}.motor.C.v", 0.0)
        x = mb.state(f"axis{i+1}.controller.PI.x", 0.0)
        # Controller (:207-216): e = ratio qd_ref + kp (ratio q_ref - phi_m) - w_m
        # observed (eliminated) variables the reference's tests read through symbolic indexing
        # (test/test_robot.jl:320-325): axisControlBus.angle_ref, .angle = motor angle / ratio
        # (robot_components.jl:134,181), and their difference
        mb.observed[f"pathPlanning.controlBus.axisControlBus{i+1}.angle_ref"] = q_ref.n
        ang = phim / ratio if typ == 1 else phi[i]
        mb.observed[f"pathPlanning.controlBus.axisControlBus{i+1}.angle"] = _node(mb, ang)
        mb.observed[f"axis{i+1}.control_error"] = _node(mb, q_ref - ang)
        e = ratio * qd_ref + kp[i] * (ratio * q_ref - phim) - wmot
        i_ref = ks[i] * (x + e)                                      # Blocks.PI: y = k (x + e)
        mb.set_rhs(x, e / Ts[i])                                     # D(x) = e / T
        # Motor (:336-408), reduced: SURVEY Appendix B.3, re-derived in DESIGN.md
        mb.set_rhs(vC, (cur - i_ref) / (10.0 * C))
        mb.set_rhs(cur, (-5.0 * vC - 250.0 * cur - k * wmot) / La)
        if typ == 1:
            # SpringDamper between rotor and gear input (:165-176): tau = c phi_rel + d w_rel
            tau_sd = c * (ratio * phi[i] - phim) + cd * (ratio * w[i] - wmot)
            mb.set_rhs(phim, wmot)
            mb.set_rhs(wmot, (k * cur + tau_sd) / J)
            joints[i].joint["tau"] = -ratio * tau_sd
        else:
            # rigid gear: joint torque = ratio (k i - J ratio qdd)
            joints[i].joint["tau"] = ratio * (k * cur - J * ratio * joints[i].acc)
    for c in range(3):                                               # tool flange position in the world
        mb.observed[f"mechanics.b6.frame_b.r_0[{c+1}]"] = _node(mb, fr.r0[c])
    mech.solve_accelerations()
    import os
    # staged temporaries (codegen/emit.py): tuning knobs of the emitted kernels, not of the model.
    # Measured on B200 (profiles/r01_robot_v5_summary.md): parking long-lived temporaries in shared
    # memory removes 45 % of the local accesses of rhs() but runs 12 % slower (volatile LDS in the
    # dependency chains, 2 warps per scheduler), so it is off (0 slots) by default.
    return mb.finish(tspan=(0.0, 2.0), mechanism=mech,
                     smem_slots=int(os.environ.get("MB_ROBOT_SLOTS", "0")),
                     strided_rhs=True,
                     reg_target=int(os.environ.get("MB_ROBOT_RTARGET", "72")),
                     stage_factor=os.environ.get("MB_ROBOT_STAGE_FACTOR", "1") == "1")


def _node(mb, e):
    return e.n if hasattr(e, "n") else mb.const(float(e)).n

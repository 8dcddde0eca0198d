"""Model zoo: the BASELINE.json configs as reduced systems.

Each builder restates one reference example with the reference's components
and default data, and returns the `ReducedSystem` that `multibody(model)` +
`ODEProblem` would hand to the integrator (src/Multibody.jl:166-188).
"""
from __future__ import annotations

from typing import Callable, Dict

import numpy as np

from .builder import ModelBuilder, ReducedSystem
from .mechanism import Mechanism
from .robot import robot6dof
from .loops import LoopMechanism


def pendulum() -> ReducedSystem:
    """C1. `World -> Revolute(n=[0,0,1], isroot) -> Body(m=1, r_cm=[0.5,0,0])`
    (docs/src/examples/pendulum.md:25-46; test/basic_tests.jl:71-119).
    States [joint.phi, joint.w] (pendulum.md:53).  The optional joint damper of
    the same page (`Rotational.Damper(d)` on the axis flange, pendulum.md:95-110)
    is included with default d = 0."""
    mb = ModelBuilder("pendulum")
    mech = Mechanism(mb)
    d = mb.param("damper.d", 0.0)
    m = mb.param("body.m", 1.0)
    phi = mb.state("joint.phi", 0.0)
    w = mb.state("joint.w", 0.0)
    jb = mech.revolute(mech.world(), (0, 0, 1), phi, w, tau_applied=-d * w, name="joint")
    mech.body(jb, m, (0.5, 0.0, 0.0))            # I_33 = 0.001 default (components.jl:345-351)
    tip = mech.fixed_translation(jb, (0.5, 0.0, 0.0))
    for c in range(2):
        mb.observed[f"body.r_0[{c+1}]"] = tip.r0[c].n    # centre of mass in the world (get_trans, orientation.jl:411)
    mech.solve_accelerations()
    mb.observed["damper.power"] = (d * w * w).n   # sensors.jl:159-173 sign aside; used by tests
    return mb.finish(tspan=(0.0, 10.0), mechanism=mech)


def spring_damper() -> ReducedSystem:
    """C2. docs/src/examples/spring_damper_system.md:21-61: `body1` is a free
    root body (Euler-angle states, components.jl:420-434) hung from `bar1` by
    `spring1 || damper1`; `body2` rides `Prismatic p2` (n = -y, s0 = 0.1) with
    `spring2` in parallel.  n = 14, explicit ODE."""
    mb = ModelBuilder("spring_damper")
    mech = Mechanism(mb)
    c1 = mb.param("spring1.c", 30.0)
    c2 = mb.param("spring2.c", 30.0)
    d1 = mb.param("damper1.d", 2.0)
    m1 = mb.param("body1.m", 1.0)
    m2 = mb.param("body2.m", 1.0)
    s0 = 0.1                                           # s_unstretched, both springs
    r0 = [mb.state(f"body1.r_0[{i+1}]", v) for i, v in enumerate((0.3, -0.2, 0.0))]
    v0 = [mb.state(f"body1.v_0[{i+1}]", 0.0) for i in range(3)]
    phi = [mb.state(f"body1.phi[{i+1}]", 0.0) for i in range(3)]
    phid = [mb.state(f"body1.phid[{i+1}]", 0.0) for i in range(3)]
    s = mb.state("p2.s", 0.1)
    v = mb.state("p2.v", 0.0)
    world = mech.world()
    bar1 = mech.fixed_translation(world, (0.3, 0.0, 0.0))
    bar2 = mech.fixed_translation(bar1, (0.6, 0.0, 0.0))
    b1 = mech.free_body_euler(r0, v0, phi, phid)
    mech.body(b1, m1, (0.0, 0.0, 0.0), (0.1, 0.1, 0.1, 0.0, 0.0, 0.0))
    mech.line_force(bar1, b1, lambda sl, ds: c1 * (sl - s0) + d1 * ds)   # spring1 || damper1
    p2 = mech.prismatic(bar2, (0.0, -1.0, 0.0), s, v, name="p2")
    mech.body(p2, m2, (0.0, -0.2, 0.0))
    mech.line_force(bar2, p2, lambda sl, ds: c2 * (sl - s0))             # spring2
    mech.solve_accelerations()
    return mb.finish(tspan=(0.0, 5.0), mechanism=mech)


def three_springs() -> ReducedSystem:
    """C2, second model.  docs/src/examples/three_springs.md:22-44 (test/basic_tests.jl:440-481):
    `body1` is a free root body with the non-unit quaternion (13 states) hanging on `spring2`
    from the massless node where `spring1` (from bar1 = [0.3,0,0]) and `spring3` (from bar2 =
    [0,0,0.3]) meet.  The node position is algebraic (force balance) -> index-1 DAE, n = 16,
    which is why the reference solves it with Rodas5P / Rodas4 and not Tsit5."""
    mb = ModelBuilder("three_springs")
    mech = Mechanism(mb)
    c1 = mb.param("spring1.c", 20.0)
    c2 = mb.param("spring2.c", 40.0)
    c3 = mb.param("spring3.c", 20.0)
    m = mb.param("body1.m", 0.8)
    s0 = 0.1
    r0 = [mb.state(f"body1.r_0[{i+1}]", v) for i, v in enumerate((0.5, -0.3, 0.0))]
    v0 = [mb.state(f"body1.v_0[{i+1}]", 0.0) for i in range(3)]
    Qh = [mb.state(f"body1.Q\u0302[{i+1}]", v) for i, v in enumerate((1.0, 0.0, 0.0, 0.0))]
    w = [mb.state(f"body1.w_a[{i+1}]", 0.0) for i in range(3)]
    rj = [mb.state(f"spring1.frame_b.r_0[{i+1}]", v, differential=False) for i, v in enumerate((0.1, -0.2, 0.2))]
    world = mech.world()
    bar1 = mech.fixed_translation(world, (0.3, 0.0, 0.0))
    bar2 = mech.fixed_translation(world, (0.0, 0.0, 0.3))
    b1 = mech.free_body_quat(r0, v0, Qh, w)
    mech.body(b1, m, (0.0, -0.2, 0.0), (0.1, 0.1, 0.1, 0.0, 0.0, 0.0))
    node = mech.junction(rj)
    mech.line_force(bar1, node, lambda sl, ds: c1 * (sl - s0))       # spring1
    mech.line_force(bar2, node, lambda sl, ds: c3 * (sl - s0))       # spring3
    mech.line_force(node, b1, lambda sl, ds: c2 * (sl - s0))         # spring2
    mech.solve_accelerations()
    for c in range(3):
        mb.observed[f"body1.r_0[{c+1}]"] = r0[c].n
    return mb.finish(tspan=(0.0, 10.0), mechanism=mech)


def free_couple() -> ReducedSystem:
    """test/test_worldforces.jl:96-213 (TestWorldForce3/4/5/7): a free `BodyShape` / `BodyCylinder`
    (r = [1,0,0], root with Euler angles, no gravity) with `WorldForce` f on frame_a and -f on
    frame_b, resolved in the world or in frame_b.  The reference prints 16-digit end positions
    of `body.frame_a.r_0` for these; they are the tightest reference-run outputs the oracle can
    be checked against (tests/golden/reference_known_answers.json)."""
    mb = ModelBuilder("free_couple")
    mech = Mechanism(mb, g=0.0)
    f = [mb.param(f"f[{i+1}]", v) for i, v in enumerate((0.0, 1.0, 0.0))]
    rw = mb.param("resolve_world", 1.0)
    m = mb.param("body.m", 1.0)
    I = [mb.param(f"body.I_{k}{k}", 0.001) for k in (1, 2, 3)]
    r0 = [mb.state(f"body.r_0[{i+1}]", 0.0) for i in range(3)]
    v0 = [mb.state(f"body.v_0[{i+1}]", 0.0) for i in range(3)]
    phi = [mb.state(f"body.phi[{i+1}]", 0.0) for i in range(3)]
    phid = [mb.state(f"body.phid[{i+1}]", 0.0) for i in range(3)]
    a = mech.free_body_euler(r0, v0, phi, phid)
    b = mech.body_shape(a, (1.0, 0.0, 0.0), m, I=(I[0], I[1], I[2], 0.0, 0.0, 0.0))
    mech.world_force(a, f, rw)
    mech.world_force(b, [-x for x in f], rw)
    mech.solve_accelerations()
    for c in range(3):
        mb.observed[f"body.frame_a.r_0[{c+1}]"] = r0[c].n
    return mb.finish(tspan=(0.0, 1.0), mechanism=mech)


def furuta() -> ReducedSystem:
    """Furuta pendulum of test/test_orientation_getters.jl:3-47 (docs/src/examples/... rotations):
    shoulder Revolute(n=y) -> upper_arm BodyShape(r=[0,0,0.6]) -> elbow Revolute(n=z, phi0=0.1) ->
    lower_arm BodyShape(r=[0,0.6,0]) -> tip Body(m=0.3); Rotational.Damper(d=0.07) on both axes;
    g = 9.81.  The reference prints `get_trans` of two frames from a Rodas5P() solve: the only
    reference-run output of the Rosenbrock path (tests/golden/reference_known_answers.json)."""
    mb = ModelBuilder("furuta")
    mech = Mechanism(mb, g=9.81)
    d1 = mb.param("damper1.d", 0.07)
    d2 = mb.param("damper2.d", 0.07)
    p1 = mb.state("shoulder_joint.phi", 0.0)
    w1 = mb.state("shoulder_joint.w", 0.0)
    p2 = mb.state("elbow_joint.phi", 0.1)
    w2 = mb.state("elbow_joint.w", 0.0)
    sh = mech.revolute(mech.world(), (0, 1, 0), p1, w1, tau_applied=-d1 * w1, name="shoulder_joint")
    ua_b = mech.body_shape(sh, (0.0, 0.0, 0.6), 1.0)
    el = mech.revolute(ua_b, (0, 0, 1), p2, w2, tau_applied=-d2 * w2, name="elbow_joint")
    la_b = mech.body_shape(el, (0.0, 0.6, 0.0), 1.0)
    mech.body(la_b, 0.3)
    for c in range(3):
        mb.observed[f"upper_arm.frame_b.r_0[{c+1}]"] = _node(mb, ua_b.r0[c])
    for c in range(3):
        mb.observed[f"lower_arm.frame_b.r_0[{c+1}]"] = _node(mb, la_b.r0[c])
    mech.solve_accelerations()
    return mb.finish(tspan=(0.0, 12.0), mechanism=mech)


def _node(mb, v):
    from .mechanism import _ex
    return _ex(mb, v).n


def planar_pendulum() -> ReducedSystem:
    """test/test_PlanarMechanics.jl:38-61: `Pl.Fixed -> Pl.Revolute -> Pl.FixedTranslation(r=[1,0]) ->
    Pl.Body(m=1, I=0.1)`, gy = -9.80665; two states (test asserts `length(unknowns) == 2`).
    The same mechanism in 3-D: Revolute about z, point mass at [1,0,0] with I_33 = 0.1.  The
    reference prints `rod.frame_a.phi` at t = 1 from a `Rodas5P()` run at default tolerances."""
    mb = ModelBuilder("planar_pendulum")
    mech = Mechanism(mb)
    phi = mb.state("body.phi", 0.0)
    w = mb.state("body.w", 0.0)
    jb = mech.revolute(mech.world(), (0, 0, 1), phi, w, name="revolute")
    mech.body(jb, 1.0, (1.0, 0.0, 0.0), (0.001, 0.001, 0.1, 0.0, 0.0, 0.0))
    mech.solve_accelerations()
    mb.observed["rod.frame_a.phi"] = phi.n
    return mb.finish(tspan=(0.0, 3.0), mechanism=mech)


def consistent_positions(sysm_g, closure, q_nodes, q0, dep, iters: int = 50, tol: float = 1e-14):
    """Newton on the closure conditions g(q) = 0 for the dependent coordinates (the
    reference leaves this to MTK's initialisation problem, test/test_fourbar.jl:44,70)."""
    g = sysm_g
    q = np.array(q0, dtype=float)
    J = g.diff(closure, [q_nodes[k] for k in dep])
    jn = [[J[i].get(j, g.zero) for j in range(len(dep))] for i in range(len(closure))]
    flat = [n for row in jn for n in row]
    for _ in range(iters):
        inp = {g.payload[qn]: float(v) for qn, v in zip(q_nodes, q)}
        vals = g.evaluate(list(closure) + flat, inp, None)
        r = np.array(vals[:len(closure)])
        A = np.array(vals[len(closure):]).reshape(len(closure), len(dep))
        if np.abs(r).max() < tol:
            break
        q[dep] -= np.linalg.solve(A, r)
    else:
        raise RuntimeError("closure Newton iteration did not converge")
    return q


def fourbar() -> ReducedSystem:
    """C3. Spatial four-bar with a prismatic leg and a cut revolute joint
    (test/test_fourbar.jl:13-42; docs/src/examples/kinematic_loops.md:90-150).
    Branch A: world -> j1 (x) -> b1 -> j3 (x) -> j4 (y) -> [j5 cut, z];
    branch B: world -> b0 [1.2,0,0] -> j2 Prismatic (x) -> b2 -> rev (y) -> rev1 (z) -> b3 -> [j5].
    States: j1.phi, j1.w differential (state_priority = 10 in the reference); the five other
    tree coordinates algebraic.  n = 7 (the reference reports 8: MTK keeps one more
    algebraic unknown; the cut angle j5.phi is eliminated here)."""
    mb = ModelBuilder("fourbar")
    mech = LoopMechanism(mb)
    phi1 = mb.state("j1.phi", 0.0)
    w1 = mb.state("j1.w", 5.235987755982989)
    s2 = mb.state("j2.s", -0.2, differential=False)
    q_rev = mb.state("rev.phi", 0.0, differential=False)
    q_rev1 = mb.state("rev1.phi", 0.0, differential=False)
    q3 = mb.state("j3.phi", 0.0, differential=False)
    q4 = mb.state("j4.phi", 0.0, differential=False)
    deps = [s2, q_rev, q_rev1, q3, q4]
    vd = {q.n: mech.dependent_velocity(q) for q in deps}
    world = mech.world()
    # joints are created in the order of q_all below
    j1 = mech.revolute(world, (1, 0, 0), phi1, w1, name="j1")
    b0 = mech.fixed_translation(world, (1.2, 0.0, 0.0))
    j2 = mech.prismatic(b0, (1, 0, 0), s2, vd[s2.n], name="j2")
    b2 = mech.body_shape(j2, (0.0, 0.2, 0.0), 1.0)
    rev = mech.revolute(b2, (0, 1, 0), q_rev, vd[q_rev.n], name="rev")
    rev1 = mech.revolute(rev, (0, 0, 1), q_rev1, vd[q_rev1.n], name="rev1")
    b3 = mech.body_shape(rev1, (-1.0, 0.3, 0.1), 1.0)
    b1 = mech.body_shape(j1, (0.0, 0.5, 0.1), 1.0)
    j3 = mech.revolute(b1, (1, 0, 0), q3, vd[q3.n], name="j3")
    j4 = mech.revolute(j3, (0, 1, 0), q4, vd[q4.n], name="j4")
    mech.cut_revolute(j4, b3, (0, 0, 1))
    q_all = [phi1, s2, q_rev, q_rev1, q3, q4]
    v_all = [w1] + [vd[q.n] for q in deps]
    mech.solve_loops(q_all, v_all)
    q0 = consistent_positions(mb.g, mech.closure, [q.n for q in q_all], [mb.u0[mb.index_of(q)] for q in q_all],
                              mech.dep)
    for q, v in zip(q_all, q0):
        mb.u0[mb.index_of(q)] = float(v)
    sysm = mb.finish(tspan=(0.0, 1.4399), mechanism=mech)
    sysm.meta["pivots"] = static_pivots(sysm)
    return sysm


def static_pivots(sysm: ReducedSystem, h: float = 1e-3):
    """Pivot sequence of the generated sparse LU of W = M/(gamma h) - J for systems whose
    algebraic rows have no usable diagonal: threshold-Markowitz on a numeric sample at u0
    (Jacobian by central differences of the interpreted RHS)."""
    from ..codegen.emit import choose_pivots_numeric
    n = sysm.n
    J = np.zeros((n, n))
    for j in range(n):
        e = np.zeros(n)
        e[j] = 1e-6
        J[:, j] = (sysm.eval_rhs(sysm.u0 + e, sysm.p0, 0.0) - sysm.eval_rhs(sysm.u0 - e, sysm.p0, 0.0)) / 2e-6
    W = np.diag(sysm.mass_diag.astype(float)) / (0.21193756319429014 * h) - J
    W[np.abs(W) < 1e-8 * np.abs(W).max()] = 0.0
    return choose_pivots_numeric(W)


def rope(n_links: int = 6, name: str = "rope6") -> ReducedSystem:
    """C5 (first milestone: the documented size).  Stiff `Rope(l=1, m=6, n=6, d_joint=1,
    dir=[0.05,1,0])` hung from the world with a `Body` at its end
    (docs/src/examples/ropes_and_cables.md:24-36; Rope: src/components.jl:574-626):
    n+1 `Spherical(state=true, d=d_joint)` joints (Euler angles, sequence [1,2,3],
    src/joints.jl:181-259) and n `BodyShape(m = m/n, r = l/n dir)` links.
    States: 6 (n+1) = 42 for n = 6."""
    mb = ModelBuilder(name)
    mech = Mechanism(mb)
    d_joint = mb.param("rope.d_joint", 1.0)
    m_body = mb.param("body.m", 1.0)
    air = mb.param("rope.air_resistance", 0.0)
    l, m_rope = 1.0, 6.0
    dirv = np.array([0.05, 1.0, 0.0])
    dirv = dirv / np.linalg.norm(dirv)
    li, mi = l / n_links, m_rope / n_links
    fr = mech.world()
    for i in range(n_links + 1):
        phi = [mb.state(f"rope.joint_{i+1}.phi[{k+1}]", 0.0) for k in range(3)]
        phid = [mb.state(f"rope.joint_{i+1}.phid[{k+1}]", 0.0) for k in range(3)]
        fr = mech.spherical(fr, phi, phid, d=d_joint)
        if i < n_links:
            fr = mech.body_shape(fr, tuple(li * dirv), mi, air_resistance=air)
    mech.body(fr, m_body, (0.0, 0.0, 0.0))
    for c in range(3):
        mb.observed[f"body.r_0[{c+1}]"] = fr.r0[c].n      # rope end (what the docs animation shows)
    mech.solve_accelerations()
    sysm = mb.finish(tspan=(0.0, 5.0), mechanism=mech)
    sysm.meta["strided_rhs"] = True        # rhs_s for the shared-memory rows of the Rosenbrock row kernel (csrc/mb_model_rope6.cu)
    return sysm


def rope_elastic(n_links: int = 3, name: str = "rope3e") -> ReducedSystem:
    """Flexible `Rope(l=1, m=5, n=3, c=800, d=0.001, d_joint=0.1, air_resistance=0.2)` with a 300 kg
    `Body` at its end (test/basic_tests.jl:928-959; docs/src/examples/ropes_and_cables.md:48-60).
    With c > 0 every link is a `Prismatic(n=dir, s0=l/n)` in parallel with `Translational.Spring(
    c n, s_rel0 = l/n)` and `Damper(d n)`, its mass a `Body(m/n, r_cm = l/(2n) dir)` on the link's
    frame_a (src/components.jl:594-611).  States 6 (n+1) + 2 n = 8 n + 6 = 30 for n = 3, the count
    the reference's test perturbs (basic_tests.jl:950-959)."""
    mb = ModelBuilder(name)
    mech = Mechanism(mb)
    d_joint = mb.param("rope.d_joint", 0.1)
    m_body = mb.param("body.m", 300.0)
    air = mb.param("rope.air_resistance", 0.2)
    c = mb.param("rope.c", 800.0)
    d = mb.param("rope.d", 0.001)
    l, m_rope = 1.0, 5.0
    dirv = np.array([0.2, 1.0, 0.0])             # released pointing upwards, as in ropes_and_cables.md:50
    dirv = dirv / np.linalg.norm(dirv)
    li, mi = l / n_links, m_rope / n_links
    ci, di = n_links * c, n_links * d
    fr = mech.world()
    for i in range(n_links + 1):
        phi = [mb.state(f"rope.joint_{i+1}.phi[{k+1}]", 0.0) for k in range(3)]
        phid = [mb.state(f"rope.joint_{i+1}.phid[{k+1}]", 0.0) for k in range(3)]
        fr = mech.spherical(fr, phi, phid, d=d_joint)
        if i < n_links:
            mech.body(fr, mi, tuple(li / 2 * dirv), air_resistance=air)
            sl = mb.state(f"rope.flexibility_{i+1}.s", li)
            vl = mb.state(f"rope.flexibility_{i+1}.v", 0.0)
            fr = mech.prismatic(fr, tuple(dirv), sl, vl, f_applied=-(ci * (sl - li) + di * vl),
                                name=f"flexibility_{i+1}")
    mech.body(fr, m_body, (0.0, 0.0, 0.0))
    for k in range(3):
        mb.observed[f"body.r_0[{k+1}]"] = fr.r0[k].n
    mech.solve_accelerations()
    sysm = mb.finish(tspan=(0.0, 5.0), mechanism=mech)
    sysm.meta["strided_rhs"] = True        # rhs_s for the shared-memory rows of the Rosenbrock row kernel
    return sysm


def fourbar_planar() -> ReducedSystem:
    """Planar loop of docs/src/examples/kinematic_loops.md:38-80: four bars of length 1.5, five
    revolute joints about z (two of them at the world origin), the loop closed by
    `RevolutePlanarLoopConstraint` j4, dampers (d = 0.1) on j1 and j2.  Two degrees of freedom:
    j1, j2 keep (phi, w) as differential states; j3 and j5 are algebraic.  j5 is traversed from
    the world side, so its coordinate here is the reversed angle -j5.phi."""
    mb = ModelBuilder("fourbar_planar")
    mech = LoopMechanism(mb)
    d1 = mb.param("damper1.d", 0.1)
    d2 = mb.param("damper2.d", 0.1)
    L = 1.5
    q1 = mb.state("j1.phi", 0.1)
    w1 = mb.state("j1.w", 0.0)
    q2 = mb.state("j2.phi", 0.0)
    w2 = mb.state("j2.w", 0.0)
    q3 = mb.state("j3.phi", 0.0, differential=False)
    q5 = mb.state("j5.phi_rev", 0.0, differential=False)
    v3, v5 = mech.dependent_velocity(q3), mech.dependent_velocity(q5)
    world = mech.world()
    j1 = mech.revolute(world, (0, 0, 1), q1, w1, tau_applied=-d1 * w1, name="j1")
    b1 = mech.body_shape(j1, (L, 0.0, 0.0), 1.0)
    j2 = mech.revolute(b1, (0, 0, 1), q2, w2, tau_applied=-d2 * w2, name="j2")
    b2 = mech.body_shape(j2, (0.0, L, 0.0), 1.0)
    j3 = mech.revolute(b2, (0, 0, 1), q3, v3, name="j3")
    b3 = mech.body_shape(j3, (-L, 0.0, 0.0), 1.0)
    j5 = mech.revolute(world, (0, 0, 1), q5, v5, name="j5")            # frame = j5.frame_a = b4.frame_b
    b4a = mech.fixed_translation(j5, (0.0, L, 0.0))                     # back along b4 (r = [0,-L,0]) to b4.frame_a
    mech.body(b4a, 1.0, (0.0, -L / 2, 0.0))
    mech.cut_revolute_planar(b3, b4a, (0, 0, 1))
    q_all = [q1, q2, q3, q5]
    v_all = [w1, w2, v3, v5]
    mech.solve_loops(q_all, v_all)
    q0 = consistent_positions(mb.g, mech.closure, [q.n for q in q_all], [mb.u0[mb.index_of(q)] for q in q_all],
                              mech.dep)
    for q, v in zip(q_all, q0):
        mb.u0[mb.index_of(q)] = float(v)
    sysm = mb.finish(tspan=(0.0, 10.0), mechanism=mech)
    sysm.meta["pivots"] = static_pivots(sysm)
    return sysm


def point_gravity_two_bodies() -> ReducedSystem:
    """test/basic_tests.jl:124-168: two free root bodies (Euler-angle roots, m = 1, I = 0.1) in the point gravity field
    of `World(point_gravity = true, mu = 1)` (src/components.jl:89-98); the reference asserts body2.r_0(5)."""
    mb = ModelBuilder("point_gravity")
    mu = mb.param("world.mu", 1.0)
    mech = Mechanism(mb, point_gravity=True, mu=mu)
    for k, (r, v) in enumerate((((0.0, 0.6, 0.0), (1.0, 0.0, 0.0)), ((0.6, 0.6, 0.0), (0.6, 0.0, 0.0)))):
        nm = f"body{k+1}"
        r0 = [mb.state(f"{nm}.r_0[{i+1}]", x) for i, x in enumerate(r)]
        v0 = [mb.state(f"{nm}.v_0[{i+1}]", x) for i, x in enumerate(v)]
        phi = [mb.state(f"{nm}.phi[{i+1}]", 0.0) for i in range(3)]
        phid = [mb.state(f"{nm}.phid[{i+1}]", 0.0) for i in range(3)]
        b = mech.free_body_euler(r0, v0, phi, phid)
        mech.body(b, 1.0, (0.0, 0.0, 0.0), (0.1, 0.1, 0.1, 0.0, 0.0, 0.0))
    mech.solve_accelerations()
    return mb.finish(tspan=(0.0, 5.0), mechanism=mech)


def spring_with_mass() -> ReducedSystem:
    """`Spring(c, m)` = `LineForceWithMass` (src/forces.jl:279-326, used by docs/src/examples/swing.md:27): a body on a
    vertical `Prismatic` hangs from the world by a spring that carries a point mass at its middle (lengthfraction 1/2).
    Closed form (Lagrange, the point mass moves with half the joint speed):
    (M + m_s / 4) s'' = M g + m_s g / 2 - c (s - s0)."""
    mb = ModelBuilder("spring_with_mass")
    mech = Mechanism(mb)
    c = mb.param("spring.c", 40.0)
    ms = mb.param("spring.m", 0.5)
    M = mb.param("body.m", 1.0)
    s = mb.state("joint.s", 0.3)
    v = mb.state("joint.v", 0.0)
    w = mech.world()
    pj = mech.prismatic(w, (0.0, -1.0, 0.0), s, v, name="joint")
    mech.body(pj, M, (0.0, 0.0, 0.0))
    mech.line_force(w, pj, lambda sl, ds: c * (sl - 0.1), m=ms, lengthfraction=0.5)
    mech.solve_accelerations()
    return mb.finish(tspan=(0.0, 5.0), mechanism=mech)


def spherical_pendulum(quat: bool) -> ReducedSystem:
    """test/basic_tests.jl:1170-1227: `World -> Spherical(state=true, quat) -> FixedTranslation(r=[1,0,0]) -> Body(m=1)`;
    the reference asserts that the quaternion and the Euler-angle joint give the same body.r_0 on 0:0.1:1 (rtol 1e-5)."""
    mb = ModelBuilder("spherical_pendulum_quat" if quat else "spherical_pendulum_euler")
    mech = Mechanism(mb)
    if quat:
        Qh = [mb.state(f"joint.Qhat[{i+1}]", 1.0 if i == 0 else 0.0) for i in range(4)]
        w = [mb.state(f"joint.w_rel[{i+1}]", 0.0) for i in range(3)]
        jb = mech.spherical_quat(mech.world(), Qh, w)
    else:
        phi = [mb.state(f"joint.phi[{i+1}]", 0.0) for i in range(3)]
        phid = [mb.state(f"joint.phid[{i+1}]", 0.0) for i in range(3)]
        jb = mech.spherical(mech.world(), phi, phid)
    fr = mech.fixed_translation(jb, (1.0, 0.0, 0.0))
    mech.body(fr, 1.0, (0.0, 0.0, 0.0))
    for c in range(3):
        v = fr.r0[c]
        mb.observed[f"body.r_0[{c+1}]"] = v.n if hasattr(v, "n") else mb.const(float(v)).n
    mech.solve_accelerations()
    return mb.finish(tspan=(0.0, 1.0), mechanism=mech)


# models that are not compiled into the library: they reach the device through an MBIR2 image that carries them
# (codegen/mbir.py; compiled at load time)
IMAGE_MODELS: Dict[str, Callable[[], ReducedSystem]] = {"point_gravity": point_gravity_two_bodies, "spring_with_mass": spring_with_mass,
                                                        "spherical_pendulum_quat": lambda: spherical_pendulum(True),
                                                        "spherical_pendulum_euler": lambda: spherical_pendulum(False)}

MODELS: Dict[str, Callable[[], ReducedSystem]] = {
    "pendulum": pendulum,
    "spring_damper": spring_damper,
    "robot6dof": robot6dof,
    "fourbar": fourbar,
    "three_springs": three_springs,
    "fourbar_planar": fourbar_planar,
    "rope6": rope,
    "free_couple": free_couple,
    "furuta": furuta,
    "planar_pendulum": planar_pendulum,
    "rope3e": rope_elastic,
}
_cache: Dict[str, ReducedSystem] = {}


def get_model(name: str) -> ReducedSystem:
    if name not in _cache:
        from .chain import CHAIN_MODELS        # loop-based long ropes (not on the expression DAG)
        _cache[name] = (MODELS[name] if name in MODELS else IMAGE_MODELS[name] if name in IMAGE_MODELS else CHAIN_MODELS[name])()
    return _cache[name]

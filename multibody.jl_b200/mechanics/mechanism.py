"""Tree-structured multibody mechanics on the expression DAG.

This module plays the part of ModelingToolkit's connection expansion + index
reduction for the component set the BASELINE configs use.  Components mirror
the reference's (`World`, `Fixed`, `FixedTranslation`, `Body`, `BodyShape`,
`Revolute`, `Prismatic`, `Spherical`, line forces), but are assembled causally
along the kinematic tree: each component maps the kinematics of `frame_a` to
`frame_b`, every frame accumulates the wrench (f, tau, resolved in the frame,
reference `Frame` src/frames.jl:1-26) its downstream side needs, and each joint
contributes one scalar equation per degree of freedom.  The joint equations
are linear in the generalised accelerations; `Mechanism.solve_accelerations`
extracts the (symmetric) mass matrix by differentiating them and inlines an
LDL^T solve into the RHS, which is what `mtkcompile(...; inline_linear_sccs =
true)` does in the reference (src/Multibody.jl:159-166,184).
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

from ..codegen.dag import Ex, sqrt, fmax
from ..codegen.emit import ldl_solve_factory
from .builder import ModelBuilder
from .orientation import (vadd, vsub, vscale, vneg, dot, cross, matvec, matTvec,
                          matmul, eye, planar_rotation, axes_rotations, from_Q)


class Frame:
    """Kinematic state of a frame: r_0, v_0, a_0 in world coordinates, R
    (world->local), w and z = dw/dt resolved locally; plus the wrench sink."""

    def __init__(self, r0, R, w, v0, a0, z, vl=None, al=None):
        self.r0, self.R, self.w, self.v0, self.a0, self.z = r0, R, w, v0, a0, z
        # the same origin velocity / (acceleration - gravity) resolved in the frame itself:
        # vl = R v0, al = R (a0 - g).  Bodies use these, so a serial chain never needs the
        # absolute rotation matrix (local-frame recursive Newton-Euler); the world-frame
        # fields above are only emitted when something references them (line forces,
        # loop closure, observed positions).
        self.vl, self.al = vl, al
        self.f = [0.0, 0.0, 0.0]     # force the downstream side needs at this frame (local)
        self.tau = [0.0, 0.0, 0.0]

    def add_wrench(self, f, tau):
        self.f = vadd(self.f, f)
        self.tau = vadd(self.tau, tau)


class Mechanism:
    def __init__(self, mb: ModelBuilder, g: float = 9.80665, n=(0.0, -1.0, 0.0), point_gravity: bool = False, mu=3.986004418e14):
        """World defaults: src/components.jl:63 (g = 9.80665, n = [0,-1,0]).  `point_gravity=True` selects the field
        -mu r / |r|^3 of `gravity_acceleration` (src/components.jl:89-98), evaluated at every body's centre of mass."""
        self.mb = mb
        self.point_gravity, self.mu = bool(point_gravity), mu
        self.gvec = [0.0, 0.0, 0.0] if point_gravity else [g * n[0], g * n[1], g * n[2]]
        self._back: List[Callable[[], None]] = []   # executed in reverse order
        self._eqs: List[tuple] = []                 # (accel Ex, velocity-state Ex, residual thunk)
        self._bodies: List[Callable[[], None]] = []
        self._junctions: List[tuple] = []           # (frame, [3 algebraic position states])
        self.frames: Dict[str, Frame] = {}

    # -- roots ---------------------------------------------------------------
    def world(self) -> Frame:
        """World / Fixed at the origin: r_b = 0, R_b = I (components.jl:76-79)."""
        z3 = [0.0, 0.0, 0.0]
        return Frame(list(z3), eye(), list(z3), list(z3), list(z3), list(z3), list(z3), vneg(self.gvec))

    def fixed(self, r) -> Frame:
        """Fixed(r): components.jl:112-113."""
        z3 = [0.0, 0.0, 0.0]
        return Frame([float(x) for x in r], eye(), list(z3), list(z3), list(z3), list(z3), list(z3),
                     vneg(self.gvec))

    # -- rigid connections ----------------------------------------------------
    def fixed_translation(self, a: Frame, r) -> Frame:
        """FixedTranslation (components.jl:247-250): r_b = r_a + R_a' r, R_b = R_a;
        f_a = f_b, tau_a = tau_b + r x f_b (wrench needed downstream)."""
        r = list(r)
        wr = cross(a.w, r)
        rel_a = vadd(cross(a.z, r), cross(a.w, wr))
        b = Frame(vadd(a.r0, matTvec(a.R, r)), a.R, a.w,
                  vadd(a.v0, matTvec(a.R, wr)),
                  vadd(a.a0, matTvec(a.R, rel_a)),
                  a.z, vadd(a.vl, wr), vadd(a.al, rel_a))

        def back():
            a.add_wrench(b.f, vadd(b.tau, cross(r, b.f)))
        self._back.append(back)
        return b

    # -- joints -----------------------------------------------------------------
    def revolute(self, a: Frame, n, phi: Ex, w: Ex, tau_applied=None, name="rev") -> Frame:
        """Revolute (joints.jl:54-70): R_b = Rrel(n,phi) R_a, w_b = Rrel w_a + n w,
        r_b = r_a, D(phi) = w, tau = -n . tau_b(flow) i.e. n . tau_needed = tau_applied."""
        mb = self.mb
        n = [float(x) for x in n]
        acc = mb.accel()
        Rrel = planar_rotation(n, phi)
        wrel = vscale(w, n)
        Rw = matvec(Rrel, a.w)
        wb = vadd(Rw, wrel)
        zb = vadd(vsub(matvec(Rrel, a.z), cross(wrel, Rw)), vscale(acc, n))
        b = Frame(a.r0, matmul(Rrel, a.R), wb, a.v0, a.a0, zb, matvec(Rrel, a.vl), matvec(Rrel, a.al))

        def back():
            a.add_wrench(matTvec(Rrel, b.f), matTvec(Rrel, b.tau))
        self._back.append(back)
        holder = {"tau": tau_applied}
        self._eqs.append((acc, w, lambda: dot(n, b.tau) - _ex(mb, holder["tau"]), holder))
        mb.set_rhs(phi, w)
        b.joint = holder
        b.acc = acc
        return b

    def prismatic(self, a: Frame, n, s: Ex, v: Ex, f_applied=None, name="prism") -> Frame:
        """Prismatic (joints.jl:139-151): r_b = r_a + R_a'(n s), R_b = R_a, D(s) = v,
        tau_a = tau_b + (n s) x f_b, f = n . f_b."""
        mb = self.mb
        n = [float(x) for x in n]
        acc = mb.accel()
        ns = vscale(s, n)
        nv = vscale(v, n)
        wns = cross(a.w, ns)
        rel_v = vadd(wns, nv)
        rel_a = vadd(vadd(cross(a.z, ns), cross(a.w, wns)),
                     vadd(vscale(2.0, cross(a.w, nv)), vscale(acc, n)))
        b = Frame(vadd(a.r0, matTvec(a.R, ns)), a.R, a.w,
                  vadd(a.v0, matTvec(a.R, rel_v)), vadd(a.a0, matTvec(a.R, rel_a)), a.z,
                  vadd(a.vl, rel_v), vadd(a.al, rel_a))

        def back():
            a.add_wrench(b.f, vadd(b.tau, cross(ns, b.f)))
        self._back.append(back)
        holder = {"tau": f_applied}
        self._eqs.append((acc, v, lambda: dot(n, b.f) - _ex(mb, holder["tau"]), holder))
        mb.set_rhs(s, v)
        b.joint = holder
        b.acc = acc
        return b

    def _euler_rotation(self, a: Frame, seq, phi, phid):
        """Shared by Spherical(state=true) and root bodies with Euler angles."""
        mb = self.mb
        accs = [mb.accel() for _ in range(3)]
        Rrel, wrel = axes_rotations(seq, phi, phid)
        # partial velocities: wrel = sum_i P_i phid_i   (P_i depends on phi only)
        g = mb.g
        dw = g.diff([c.n if isinstance(c, Ex) else g.const(c) for c in wrel],
                    [p.n for p in phid])
        P = [[Ex(g, dw[c].get(i, g.zero)) for c in range(3)] for i in range(3)]
        # d(wrel)/dt = sum_i P_i acc_i + sum_i (dP_i/dt) phid_i
        dphi = g.diff([c.n if isinstance(c, Ex) else g.const(c) for c in wrel],
                      [p.n for p in phi])
        wrel_dot = []
        for c in range(3):
            e = mb.const(0.0)
            for i in range(3):
                e = e + P[i][c] * accs[i] + Ex(g, dphi[c].get(i, g.zero)) * phid[i]
            wrel_dot.append(e)
        return accs, Rrel, wrel, wrel_dot, P

    def spherical(self, a: Frame, phi: Sequence[Ex], phid: Sequence[Ex], d=0.0,
                  seq=(1, 2, 3)) -> Frame:
        """Spherical(state=true) (joints.jl:213-259): Rrel = axes_rotations(seq, phi,
        phid), R_b = Rrel R_a, r_b = r_a; damping tau = -d w_rel."""
        mb = self.mb
        accs, Rrel, wrel, wrel_dot, P = self._euler_rotation(a, seq, phi, phid)
        Rw = matvec(Rrel, a.w)
        wb = vadd(Rw, wrel)
        zb = vadd(vsub(matvec(Rrel, a.z), cross(wrel, Rw)), wrel_dot)
        b = Frame(a.r0, matmul(Rrel, a.R), wb, a.v0, a.a0, zb, matvec(Rrel, a.vl), matvec(Rrel, a.al))

        def back():
            a.add_wrench(matTvec(Rrel, b.f), matTvec(Rrel, b.tau))
        self._back.append(back)
        # joints.jl:218-222 resolves the friction torque d*w_rel through Rrel
        # (`fric ~ resolve1(Rrel, frame_b.tau)`); kept as the reference has it.
        fric = matvec(Rrel, vscale(d, wrel)) if not _is_zero(d) else [0.0, 0.0, 0.0]
        for i in range(3):
            self._eqs.append((accs[i], phid[i],
                              (lambda i=i: dot(P[i], vadd(b.tau, fric))), {}))
            mb.set_rhs(phi[i], phid[i])
        return b

    def spherical_quat(self, a: Frame, Qh: Sequence[Ex], w: Sequence[Ex], d=0.0, k: float = 0.1) -> Frame:
        """Spherical(state=true, quat=true) (joints.jl:226-233): the relative orientation is the non-unit quaternion of
        orientation.jl:438-472 -- D(Qhat) = Omega(w_rel) Qhat / 2 + k (1 - Qhat'Qhat) Qhat, Rrel = from_Q(Qhat / |Qhat|) --
        and the states are Qhat[4] and w_rel[3] (resolved in frame_b); no Euler-angle singularity."""
        mb = self.mb
        n = Qh[0] * Qh[0] + Qh[1] * Qh[1] + Qh[2] * Qh[2] + Qh[3] * Qh[3]
        c = k * (1.0 - n)
        Om = [[0.0, -w[0], -w[1], -w[2]], [w[0], 0.0, w[2], -w[1]], [w[1], -w[2], 0.0, w[0]], [w[2], w[1], -w[0], 0.0]]
        for i in range(4):
            acc = c * Qh[i]
            for j in range(4):
                if not (isinstance(Om[i][j], float) and Om[i][j] == 0.0):
                    acc = acc + 0.5 * (Om[i][j] * Qh[j])
            mb.set_rhs(Qh[i], acc)
        rn = 1.0 / sqrt(n)
        Rrel = from_Q([q * rn for q in Qh])
        accs = [mb.accel() for _ in range(3)]
        wrel = list(w)
        Rw = matvec(Rrel, a.w)
        wb = vadd(Rw, wrel)
        zb = vadd(vsub(matvec(Rrel, a.z), cross(wrel, Rw)), accs)
        b = Frame(a.r0, matmul(Rrel, a.R), wb, a.v0, a.a0, zb, matvec(Rrel, a.vl), matvec(Rrel, a.al))

        def back():
            a.add_wrench(matTvec(Rrel, b.f), matTvec(Rrel, b.tau))
        self._back.append(back)
        fric = matvec(Rrel, vscale(d, wrel)) if not _is_zero(d) else [0.0, 0.0, 0.0]    # resolved as joints.jl:218-222
        for i in range(3):
            self._eqs.append((accs[i], w[i], (lambda i=i: vadd(b.tau, fric)[i]), {}))
        return b

    def free_body_euler(self, r0: Sequence[Ex], v0: Sequence[Ex], phi: Sequence[Ex],
                        phid: Sequence[Ex], seq=(1, 2, 3)) -> Frame:
        """Root `Body(isroot=true)` with Euler angles (components.jl:420-434): states
        r_0, v_0, phi, phid; R = axes_rotations(seq, phi, phid)."""
        mb = self.mb
        wf = self.world()
        accs, Rrel, wrel, wrel_dot, P = self._euler_rotation(wf, seq, phi, phid)
        acc_t = [mb.accel() for _ in range(3)]
        b = Frame(list(r0), Rrel, wrel, list(v0), list(acc_t), wrel_dot,
                  matvec(Rrel, list(v0)), matvec(Rrel, vsub(list(acc_t), self.gvec)))
        for i in range(3):
            self._eqs.append((accs[i], phid[i], (lambda i=i: dot(P[i], b.tau)), {}))
            mb.set_rhs(phi[i], phid[i])
        for i in range(3):
            self._eqs.append((acc_t[i], v0[i], (lambda i=i: matTvec(b.R, b.f)[i]), {}))
            mb.set_rhs(r0[i], v0[i])
        return b

    def free_body_quat(self, r0: Sequence[Ex], v0: Sequence[Ex], Qh: Sequence[Ex], w: Sequence[Ex],
                       k: float = 0.1) -> Frame:
        """Root `Body(isroot=true, quat=true)` (components.jl:412-418) with the non-unit
        quaternion of orientation.jl:438-472: states r_0, v_0, Qhat[4], w_a[3];
        D(Qhat) = Omega(w) Qhat / 2 + k (1 - Qhat'Qhat) Qhat,  Q = Qhat / |Qhat|,  R = from_Q(Q)."""
        mb = self.mb
        n = Qh[0] * Qh[0] + Qh[1] * Qh[1] + Qh[2] * Qh[2] + Qh[3] * Qh[3]
        c = k * (1.0 - n)
        Om = [[0.0, -w[0], -w[1], -w[2]], [w[0], 0.0, w[2], -w[1]], [w[1], -w[2], 0.0, w[0]], [w[2], w[1], -w[0], 0.0]]
        for i in range(4):
            acc = c * Qh[i]
            for j in range(4):
                if not (isinstance(Om[i][j], float) and Om[i][j] == 0.0):
                    acc = acc + 0.5 * (Om[i][j] * Qh[j])
            mb.set_rhs(Qh[i], acc)
        rn = 1.0 / sqrt(n)
        R = from_Q([q * rn for q in Qh])
        acc_r = [mb.accel() for _ in range(3)]
        acc_t = [mb.accel() for _ in range(3)]
        b = Frame(list(r0), R, list(w), list(v0), list(acc_t), list(acc_r),
                  matvec(R, list(v0)), matvec(R, vsub(list(acc_t), self.gvec)))
        for i in range(3):
            self._eqs.append((acc_r[i], w[i], (lambda i=i: b.tau[i]), {}))
        for i in range(3):
            self._eqs.append((acc_t[i], v0[i], (lambda i=i: matTvec(b.R, b.f)[i]), {}))
            mb.set_rhs(r0[i], v0[i])
        return b

    def junction(self, r: Sequence[Ex]) -> Frame:
        """Massless connection point of force elements (the node where spring1, spring2 and
        spring3 of docs/src/examples/three_springs.md:22-44 meet): its position is an algebraic
        state fixed by the force balance 0 = sum of the forces acting on it."""
        z3 = [0.0, 0.0, 0.0]
        fr = Frame(list(r), eye(), list(z3), list(z3), list(z3), list(z3), list(z3), vneg(self.gvec))
        self._junctions.append((fr, list(r)))
        return fr

    # -- bodies and forces ------------------------------------------------------
    def body(self, fr: Frame, m, r_cm=(0.0, 0.0, 0.0), I=(0.001, 0.001, 0.001, 0.0, 0.0, 0.0),
             air_resistance=0.0):
        """Body (components.jl:445-458):
        f = m (R (a_0 - g_0 [+ c |v_0| v_0]) + z x r_cm + w x (w x r_cm))
        tau = I z + w x (I w) + r_cm x f ;  I = (I11, I22, I33, I21, I31, I32)."""
        r_cm = list(r_cm)
        I11, I22, I33, I21, I31, I32 = I
        Im = [[I11, I21, I31], [I21, I22, I32], [I31, I32, I33]]

        def apply():
            # R (a_0 - g_0 + c |v_0| v_0) = al + c |vl| vl  (rotations preserve the norm)
            acc = fr.al
            if self.point_gravity:
                # Body: a_0 - gravity_acceleration(r_0 + resolve1(R, r_cm)) (src/components.jl:445-447); the base
                # acceleration trick of the uniform field does not apply, the field depends on where the body is
                rw = vadd(fr.r0, matTvec(fr.R, r_cm))
                n2 = dot(rw, rw)
                gw = vscale(-self.mu / (n2 * sqrt(n2)), rw)
                acc = vsub(acc, matvec(fr.R, gw))
            v2 = dot(fr.vl, fr.vl)
            if not _is_zero(air_resistance) and isinstance(v2, Ex):
                # |v| v; the 1e-300 keeps d|v|/dv = v/|v| finite (= 0) at rest, where the
                # reference's finite-difference Jacobian (autodiff=false) sees 0 as well
                nv = sqrt(v2 + 1e-300)
                acc = vadd(acc, vscale(air_resistance * nv, fr.vl))
            lin = vadd(acc, vadd(cross(fr.z, r_cm), cross(fr.w, cross(fr.w, r_cm))))
            f = vscale(m, lin)
            tau = vadd(vadd(matvec(Im, fr.z), cross(fr.w, matvec(Im, fr.w))), cross(r_cm, f))
            fr.add_wrench(f, tau)
        self._bodies.append(apply)

    def body_shape(self, a: Frame, r, m, r_cm=None, I=(0.001, 0.001, 0.001, 0.0, 0.0, 0.0),
                   air_resistance=0.0) -> Frame:
        """BodyShape (components.jl:509-541): Body at frame_a + FixedTranslation(r);
        default r_cm = r/2 (components.jl:475)."""
        if r_cm is None:
            r_cm = [0.5 * x for x in r]
        self.body(a, m, r_cm, I, air_resistance)
        return self.fixed_translation(a, r)

    def line_force(self, a: Frame, b: Frame, law: Callable[[Ex, Ex], Ex], m=0.0, lengthfraction: float = 0.5):
        """Line force element (forces.jl:255-258, 348-354): r_rel_0 = r_b - r_a,
        s = max(|r_rel_0|, 1e-10), e = r_rel_0/s; `law(s, ds)` is the force f
        (Spring f = c (s - s0) forces.jl:392-453; Damper f = d ds :489;
        SpringDamperParallel :524-527).  Positive f pulls the frames together.
        `m` > 0: `LineForceWithMass` (forces.jl:279-326) -- a point mass on the line at `lengthfraction` of the way from
        frame_a to frame_b: r_CM = r_a + lengthfraction r_rel, so its acceleration is the same weighting of the two frame
        accelerations; frame_a carries (1 - lengthfraction) m (a_CM - g), frame_b lengthfraction m (a_CM - g)."""
        lf = float(lengthfraction)
        has_mass = not _is_zero(m)

        def apply():
            if has_mass:
                rcm = vadd(vscale(1.0 - lf, a.r0), vscale(lf, b.r0))
                acm = vadd(vscale(1.0 - lf, a.a0), vscale(lf, b.a0))
                if self.point_gravity:
                    n2 = dot(rcm, rcm)
                    gw = vscale(-self.mu / (n2 * sqrt(n2)), rcm)
                else:
                    gw = self.gvec
                ag = vsub(acm, gw)
                a.add_wrench(matvec(a.R, vscale((1.0 - lf) * m, ag)), [0.0, 0.0, 0.0])
                b.add_wrench(matvec(b.R, vscale(lf * m, ag)), [0.0, 0.0, 0.0])
            rrel = vsub(b.r0, a.r0)
            length = sqrt(dot(rrel, rrel))
            s = fmax(length, 1e-10)
            e = [c / s for c in rrel]
            ds = dot(e, vsub(b.v0, a.v0))
            f = law(s, ds)
            fw = vscale(f, e)               # world force applied on a; -fw on b
            a.add_wrench(vneg(matvec(a.R, fw)), [0.0, 0.0, 0.0])
            b.add_wrench(matvec(b.R, fw), [0.0, 0.0, 0.0])
        self._bodies.append(apply)

    def world_force(self, b: Frame, f, resolve_world=1.0):
        """WorldForce (forces.jl:161-170, 216-226): external force on frame b, `f` given in world
        coordinates (`resolve_frame=:world`, frame_b.f ~ -resolve2(R_b, f)) or in frame_b itself
        (`:frame_b`, frame_b.f ~ -f).  `resolve_world` is 1 / 0 or a parameter blending the two,
        so one compiled kernel set serves both variants."""
        def apply():
            fl = list(f)
            fw = matvec(b.R, fl)
            if isinstance(resolve_world, (int, float)):
                F = fw if resolve_world else fl
            else:
                F = [resolve_world * fw[i] + (1.0 - resolve_world) * fl[i] for i in range(3)]
            b.add_wrench(vneg(F), [0.0, 0.0, 0.0])
        self._bodies.append(apply)

    # -- assembly -----------------------------------------------------------------
    def solve_accelerations(self):
        """Close the model: accumulate wrenches leaf->root, form the joint equations
        g(u, a) = 0, split them into M(u) a + bias(u) and inline the LDL^T solve."""
        mb = self.mb
        g = mb.g
        for apply in self._bodies:
            apply()
        for back in reversed(self._back):
            back()
        for fr, rs in self._junctions:              # algebraic rows: force balance of the node
            for i in range(3):
                mb.set_rhs(rs[i], _ex(mb, fr.f[i]))
        accs = [e[0] for e in self._eqs]
        res = [_ex(mb, e[2]()).n for e in self._eqs]
        acc_nodes = [a.n for a in accs]
        dres = g.diff(res, acc_nodes)
        k = len(accs)
        M = [[dres[i].get(j, g.zero) for j in range(k)] for i in range(k)]
        zero_map = {a: g.zero for a in acc_nodes}
        bias = g.subs(res, zero_map)
        solve = ldl_solve_factory(g, M)
        sol = solve([g.neg(b) for b in bias])
        for (acc, vel, _, _), s in zip(self._eqs, sol):
            mb.set_rhs(vel, Ex(g, s))
        # Jacobian rows of the velocity states by implicit differentiation:
        #   M a + h(u) = 0  =>  da/du_j = -M^-1 (d res/d u_j)|_{a = solution}
        # (reuses the LDL^T factors of the RHS; ~40 % fewer flops and much less code than
        # differentiating through the factorisation).
        n_state = len(mb.state_names)
        u_nodes = [g.input("u", i) for i in range(n_state)]
        dres_du = g.diff(res, u_nodes)
        amap = {a: s_ for a, s_ in zip(acc_nodes, sol)}
        rows = {mb.index_of(vel): {} for (_, vel, _, _) in self._eqs}
        vel_idx = [mb.index_of(vel) for (_, vel, _, _) in self._eqs]
        for j in range(n_state):
            col = [dres_du[i].get(j, g.zero) for i in range(k)]
            if all(c == g.zero for c in col):
                continue
            col = g.subs(col, amap)
            x = solve([g.neg(c) for c in col])
            for i in range(k):
                if x[i] != g.zero:
                    rows[vel_idx[i]][j] = x[i]
        mb.jac_rows.update(rows)
        self.residuals = res
        self.acc_nodes = acc_nodes
        self.acc_solution = sol
        self.M = M
        self._ldl = solve
        return sol


def _ex(mb: ModelBuilder, v) -> Ex:
    if v is None:
        return mb.const(0.0)
    return v if isinstance(v, Ex) else mb.const(float(v))


def _is_zero(v) -> bool:
    if isinstance(v, Ex):
        return v.is_zero()
    return float(v) == 0.0

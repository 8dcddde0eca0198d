"""Closed kinematic loops: tree mechanism + cut joint -> index-1 mass-matrix DAE.

The reference closes a loop by marking one revolute joint `iscut=true`
(src/joints.jl:25-91: `frame_a.r_0 ~ frame_b.r_0` plus the three `residue`
angles of src/orientation.jl:232-247 with the cut joint's own angle as an
extra unknown) and lets ModelingToolkit's Pantelides/dummy-derivative pass
(`mtkcompile`, src/Multibody.jl:184) reduce the index-3 system: one joint keeps
its (phi, w) as differential states, everything else becomes algebraic
(docs/src/examples/kinematic_loops.md:113,218: 2 differential + 6 algebraic).

Here the same reduction is done constructively.  With q = (q_I, q_D) the tree
joint coordinates (independent / dependent) and g(q) = 0 the five scalar
closure conditions of a cut revolute (3 position, 2 orientation; equivalent to
the reference's 6 residues with the cut angle eliminated):

  states        q_I, v_I            differential      q_I' = v_I,  v_I' = a_I
                q_D                 algebraic         0 = g(q)
  inline        v_D = S_D v_I       S_D = -G_D^-1 G_I                   (G = dg/dq)
                a   = S a_I + c     S = [I; S_D],  c = [0; -G_D^-1 gamma], gamma = d(Gv)/dq v
                a_I                 from (S' M S) a_I = -S' (M c + h)

i.e. the joint equations M a + h = B lambda of the tree are projected on the null space
of the constraints (coordinate partitioning); the ideal cut forces drop out, so the tree
may contain joints without any inertia behind them (j3, j4 of the four-bar).  These inline
solves are the "linear SCCs" MTK inlines into the RHS (`inline_linear_sccs=true`,
src/Multibody.jl:159-184).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

from ..codegen.dag import Ex
from ..codegen.emit import ldl_solve_factory
from .mechanism import Mechanism, Frame, _ex
from .orientation import vsub, vadd, vneg, vscale, matvec, matTvec, matmul, transpose, dot


def _perp(n):
    """Two unit vectors spanning the plane normal to the unit axis n."""
    import numpy as np
    n = np.asarray(n, float)
    t = np.array([1.0, 0.0, 0.0]) if abs(n[0]) < 0.9 else np.array([0.0, 1.0, 0.0])
    e1 = np.cross(n, t); e1 /= np.linalg.norm(e1)
    e2 = np.cross(n, e1)
    return [float(x) for x in e1], [float(x) for x in e2]


class LoopMechanism(Mechanism):
    def __init__(self, mb, **kw):
        super().__init__(mb, **kw)
        self._cuts: List[dict] = []
        self._dep: Dict[int, Ex] = {}        # placeholder node of a dependent velocity -> its position state

    # -- dependent joints: position is an algebraic state, velocity an inline unknown ------
    def dependent_velocity(self, q: Ex) -> Ex:
        v = self.mb.accel()
        self._dep[v.n] = q
        return v

    def cut_revolute(self, a: Frame, b: Frame, n):
        """Revolute(iscut=true) between frame a and frame b, axis n (resolved in a)."""
        mb = self.mb
        e1, e2 = _perp(n)
        cut = dict(a=a, b=b, n=[float(x) for x in n], e1=e1, e2=e2)
        self._cuts.append(cut)
        return cut

    def cut_revolute_planar(self, a: Frame, b: Frame, n):
        """RevolutePlanarLoopConstraint (src/joints.jl:593-651): for mechanisms moving in the plane
        normal to n only the two in-plane components of frame_a.r_0 ~ frame_b.r_0 are imposed
        (the joint transmits only the in-plane constraint forces f_c)."""
        e1, e2 = _perp(n)
        cut = dict(a=a, b=b, n=[float(x) for x in n], e1=e1, e2=e2, planar=True)
        self._cuts.append(cut)
        return cut

    @staticmethod
    def closure_residuals(cut) -> List[Ex]:
        a, b = cut["a"], cut["b"]
        r_rel = matvec(a.R, vsub(b.r0, a.r0))                   # frame_a.r_0 ~ frame_b.r_0
        if cut.get("planar"):
            return [dot(cut["e1"], r_rel), dot(cut["e2"], r_rel)]
        Rrel = matmul(b.R, transpose(a.R))                      # a-coordinates -> b-coordinates
        nb = matvec(Rrel, cut["n"])                             # the axis must be the same in both
        return [r_rel[0], r_rel[1], r_rel[2], dot(cut["e1"], nb), dot(cut["e2"], nb)]

    def solve_loops(self, q_all: Sequence[Ex], v_all: Sequence[Ex]):
        """q_all/v_all: every tree joint coordinate with its velocity expression (a state for
        independent joints, the `dependent_velocity` placeholder for dependent ones), in the
        order the joints were created (= order of self._eqs)."""
        mb, g = self.mb, self.mb.g
        assert len(q_all) == len(self._eqs) == len(v_all)
        # ---- position and velocity level -------------------------------------------------
        gres = [_ex(mb, e) for cut in self._cuts for e in self.closure_residuals(cut)]
        nc = len(gres)
        qn = [q.n for q in q_all]
        G = g.diff([e.n for e in gres], qn)                       # rows: {joint index: node}
        dep = [k for k, v in enumerate(v_all) if v.n in self._dep]
        ind = [k for k in range(len(q_all)) if k not in dep]
        assert len(dep) == nc, "need as many dependent coordinates as closure conditions"
        # G_D x = r via the normal equations (SPD: no pivoting needed while G_D is regular)
        A = [[G[i].get(k, g.zero) for k in dep] for i in range(nc)]
        AtA = [[self._sum(g, [g.mul(A[r][i], A[r][j]) for r in range(nc)]) for j in range(nc)] for i in range(nc)]
        nsolve = ldl_solve_factory(g, AtA)

        def solve_D(r):
            return nsolve([self._sum(g, [g.mul(A[q][i], r[q]) for q in range(nc)]) for i in range(nc)])
        nj = len(q_all)
        # S: columns per independent coordinate (identity on `ind`, S_D on `dep`)
        S = []
        for k in ind:
            sd = solve_D([g.neg(G[i].get(k, g.zero)) for i in range(nc)])
            col = [g.zero] * nj
            col[k] = g.one
            for j, kd in enumerate(dep):
                col[kd] = sd[j]
            S.append(col)
        vD = [self._sum(g, [g.mul(S[c][kd], v_all[k].n) for c, k in enumerate(ind)]) for kd in dep]
        vmap = {v_all[k].n: vD[j] for j, k in enumerate(dep)}
        # gamma = d(G v)/dq . v  with the velocities held fixed
        Gv = [self._sum(g, [g.mul(G[i].get(k, g.zero), v_all[k].n) for k in range(nj)]) for i in range(nc)]
        dGv = g.diff(Gv, qn)
        gamma = [self._sum(g, [g.mul(dGv[i].get(k, g.zero), v_all[k].n) for k in range(nj)]) for i in range(nc)]
        cD = solve_D([g.neg(x) for x in gamma])
        cvec = [g.zero] * nj
        for j, kd in enumerate(dep):
            cvec[kd] = cD[j]
        # ---- acceleration level: tree joint equations projected on the constraint null space ------
        for apply in self._bodies:
            apply()
        for back in reversed(self._back):
            back()
        accs = [e[0].n for e in self._eqs]
        res = [_ex(mb, e[2]()).n for e in self._eqs]
        d = g.diff(res, accs)
        M = [[d[i].get(j, g.zero) for j in range(nj)] for i in range(nj)]
        h = g.subs(res, {x: g.zero for x in accs})
        Mc_h = [g.add(self._sum(g, [g.mul(M[i][j], cvec[j]) for j in range(nj)]), h[i]) for i in range(nj)]
        MS = [[self._sum(g, [g.mul(M[i][j], S[c][j]) for j in range(nj)]) for i in range(nj)] for c in range(len(ind))]
        Mred = [[self._sum(g, [g.mul(S[a_][i], MS[b_][i]) for i in range(nj)]) for b_ in range(len(ind))]
                for a_ in range(len(ind))]
        rred = [g.neg(self._sum(g, [g.mul(S[a_][i], Mc_h[i]) for i in range(nj)])) for a_ in range(len(ind))]
        aI = ldl_solve_factory(g, Mred)(rred)
        a_sol = [g.add(self._sum(g, [g.mul(S[c][k], aI[c]) for c in range(len(ind))]), cvec[k]) for k in range(nj)]
        # ---- state equations -------------------------------------------------------------------------
        a_sol = g.subs(a_sol, vmap)
        for k in ind:
            mb.set_rhs(v_all[k], Ex(g, a_sol[k]))
            mb.set_rhs(q_all[k], v_all[k])
        for j, k in enumerate(dep):
            mb.set_rhs(q_all[k], gres[j])                           # algebraic row: 0 = g_j(q)
        self.closure = [e.n for e in gres]
        self.G = G
        self.dep, self.ind = dep, ind
        self.vD = vD
        self.acc_solution = a_sol
        return a_sol

    @staticmethod
    def _sum(g, terms):
        s = g.zero
        for t in terms:
            s = g.add(s, t)
        return s

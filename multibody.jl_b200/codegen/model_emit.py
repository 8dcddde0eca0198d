"""ReducedSystem -> one self-contained C / CUDA C model header.

North star (1): "a new codegen backend that takes the MTK-generated RHS, mass
matrix and Jacobian of the index-reduced system ... and emits straight-line
CUDA C compiled for sm_100a".  The reduced system is the `ReducedSystem`
built by `mechanics/` (stand-in for `mtkcompile`, reference
src/Multibody.jl:184); this module prints, per model `X`:

  mbm_X_rhs(u,p,t,tb,du)          du = f(u,p,t)          (prob.f, src/Multibody.jl:69)
  mbm_X_jac(u,p,t,tb,Jnz)         analytic sparse df/du, fixed pattern (oracle, tests)
  mbm_X_dT(u,p,t,tb,dT)           df/dt
  mbm_X_factor(u,p,t,tb,c,LU)     W = c*M - J (c = 1/(gamma*h)) built and LU-factored
                                  in one straight-line pass, pivot sequence fixed at
                                  code-generation time
  mbm_X_solve(LU,c,p,b)           in-place W^-1 b with those factors (state-independent cheap
                                  factors are recomputed from c and p instead of being stored)

plus sizes, the sparsity pattern, the mass-matrix diagonal and the algorithmic
flop counts of each function (BASELINE.md section 4 cost table).

The includer defines MB_FN (e.g. `static inline` or `__device__ __noinline__`)
and provides `mb_interp_{cubic,linear}{0,1,2,3}(tb, table, t)`.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .dag import Graph, Ex, CONST
from .emit import Printer, sparse_lu, lu_solve


class ModelCode:
    """Everything the emitter derived from a ReducedSystem (also used by tests)."""

    def __init__(self, sys, pivots: Optional[Sequence[Tuple[int, int]]] = None, recompute_limit: float = 12.0):
        self.sys = sys
        g: Graph = sys.g
        n = sys.n
        self.n, self.m = n, sys.m
        u_nodes = sys.u_nodes()
        t_node = sys.t_node()
        # --- analytic Jacobian (sparse) and time derivative -------------------
        override = getattr(sys, "jac_rows", None) or {}
        plain_rows = [i for i in range(n) if i not in override]
        Jp = g.diff([sys.rhs[i] for i in plain_rows], u_nodes)
        J = [None] * n
        for i, r in zip(plain_rows, Jp):
            J[i] = r
        for i, r in override.items():
            J[i] = {j: v for j, v in r.items() if v != g.zero}
        self.jac_entries: List[Tuple[int, int, int]] = []      # (row, col, node)
        for i in range(n):
            for j in sorted(J[i]):
                self.jac_entries.append((i, j, J[i][j]))
        dT = g.diff(sys.rhs, [t_node])
        self.dT_nodes = [dT[i].get(0, g.zero) for i in range(n)]
        # --- fused W build + LU ------------------------------------------------
        c = g.input("c", 0)
        W: Dict[Tuple[int, int], int] = {}
        for i, j, node in self.jac_entries:
            W[(i, j)] = g.neg(node)
        for i in range(n):
            if sys.mass_diag[i]:
                W[(i, i)] = g.add(W.get((i, i), g.zero), c)
        self.W_pattern = sorted(W)
        order, invd, L, U = sparse_lu(g, W, n, pivots=pivots)
        # LU storage: for each elimination step, [1/pivot, U row entries..., L column entries...].
        # Structural constants take no storage.  Entries that do not depend on the state or on
        # time (the drive blocks of the robot are linear: their factors are functions of
        # c = 1/(gamma h) and the parameters only) and are cheap are NOT stored either: the
        # solve recomputes them from (c, p, stored slots).  The kernels are bound by
        # thread-local memory traffic, not by the FP64 pipe, so a few extra flops per solve
        # are cheaper than a slot that is written once and read by all 8 stage solves.
        self.lu_nodes: List[int] = []
        self.lu_slots: Dict[Tuple[str, int, int], int] = {}
        self.lu_const: Dict[Tuple[str, int, int], float] = {}
        self.lu_recomp: Dict[Tuple[str, int, int], int] = {}
        from .dag import INPUT, INTERP, ADD, SUB, MUL, NEG
        dep_cache: Dict[int, bool] = {}

        def state_dep(n: int) -> bool:
            stack = [n]
            while stack:
                k = stack[-1]
                if k in dep_cache:
                    stack.pop(); continue
                o = g.op[k]
                if o == CONST:
                    dep_cache[k] = False; stack.pop(); continue
                if o == INPUT:
                    dep_cache[k] = g.payload[k][0] not in ("c", "p"); stack.pop(); continue
                if o == INTERP:
                    dep_cache[k] = True; stack.pop(); continue
                kids = [g.a[k]] + ([g.b[k]] if g.b[k] >= 0 else []) + ([g.c[k]] if g.c[k] >= 0 and o not in (INTERP,) else [])
                todo = [q for q in kids if q not in dep_cache]
                if todo:
                    stack.extend(todo); continue
                dep_cache[k] = any(dep_cache[q] for q in kids)
                stack.pop()
            return dep_cache[n]
        stored_node: Dict[int, int] = {}          # node -> slot (first slot holding it)
        cost_cache: Dict[int, float] = {}

        def recomp_cost(n: int) -> float:
            """ops needed to rebuild n from (c, p, constants, stored slots); inf if not allowed"""
            if n in stored_node:
                return 0.0
            if n in cost_cache:
                return cost_cache[n]
            o = g.op[n]
            if o == CONST or (o == INPUT and g.payload[n][0] in ("c", "p")):
                v = 0.0
            elif o in (ADD, SUB, MUL):
                v = 1.0 + recomp_cost(g.a[n]) + recomp_cost(g.b[n])
            elif o == NEG:
                v = recomp_cost(g.a[n])
            else:
                v = float("inf")
            cost_cache[n] = v
            return v

        def slot(kind, k, idx, node):
            if g.op[node] == CONST:
                self.lu_const[(kind, k, idx)] = g.val[node]
                return
            if recompute_limit > 0 and not state_dep(node) and recomp_cost(node) <= recompute_limit:
                self.lu_recomp[(kind, k, idx)] = node
                return
            if node in stored_node:               # same value already stored: share the slot
                self.lu_slots[(kind, k, idx)] = stored_node[node]
                return
            self.lu_slots[(kind, k, idx)] = len(self.lu_nodes)
            stored_node[node] = len(self.lu_nodes)
            cost_cache.clear()
            self.lu_nodes.append(node)
        for step, (pr, pc) in enumerate(order):
            slot("d", step, 0, invd[step])
            for j, v in U[step]:
                slot("u", step, j, v)
            for i, l in L[step]:
                slot("l", step, i, l)
        self.order, self.L, self.U, self.invd = order, L, U, invd
        # --- solve graph (separate, tiny) ------------------------------------------
        sg = Graph()
        self.solve_graph = sg
        copy_memo: Dict[int, int] = {}

        def copy(n: int) -> int:
            if n in stored_node:
                return sg.input("lu", stored_node[n])
            if n in copy_memo:
                return copy_memo[n]
            o = g.op[n]
            if o == CONST:
                r = sg.const(g.val[n])
            elif o == INPUT:
                r = sg.input(*g.payload[n])
            elif o == NEG:
                r = sg.neg(copy(g.a[n]))
            else:
                x, y = copy(g.a[n]), copy(g.b[n])
                r = {ADD: sg.add, SUB: sg.sub, MUL: sg.mul}[o](x, y)
            copy_memo[n] = r
            return r

        def lu_in(kind, k, idx):
            if (kind, k, idx) in self.lu_const:
                return sg.const(self.lu_const[(kind, k, idx)])
            if (kind, k, idx) in self.lu_recomp:
                return copy(self.lu_recomp[(kind, k, idx)])
            return sg.input("lu", self.lu_slots[(kind, k, idx)])
        sL = {k: [(i, lu_in("l", k, i)) for i, _ in L[k]] for k in range(n)}
        sU = {k: [(j, lu_in("u", k, j)) for j, _ in U[k]] for k in range(n)}
        sD = {k: lu_in("d", k, 0) for k in range(n)}
        b = [sg.input("b", i) for i in range(n)]
        self.solve_out = lu_solve(sg, order, sD, sL, sU, b)
        # --- flop counts -----------------------------------------------------------
        self.F_rhs = g.flops(sys.rhs)
        self.F_jac = g.flops([e[2] for e in self.jac_entries])
        self.F_dT = g.flops(self.dT_nodes)
        self.F_factor = g.flops(self.lu_nodes)
        self.F_solve = sg.flops(self.solve_out)
        self.F_lu_only = self.F_factor - self.F_jac   # structured LU alone (approx.; CSE shared)

    @property
    def nnz(self) -> int:
        return len(self.jac_entries)

    @property
    def nlu(self) -> int:
        return len(self.lu_nodes)


def _carray(ctype: str, name: str, vals, per_line: int = 16) -> str:
    vals = list(vals)
    if not vals:
        return f"static const {ctype} {name}[1] = {{0}};"
    rows = []
    for k in range(0, len(vals), per_line):
        rows.append("  " + ", ".join(str(v) for v in vals[k:k + per_line]))
    return f"static const {ctype} {name}[{len(vals)}] = {{\n" + ",\n".join(rows) + "\n};"


def emit_model_header(sys, code: Optional[ModelCode] = None) -> str:
    code = code or ModelCode(sys)
    g = sys.g
    X = sys.name
    n, m = sys.n, sys.m
    fmt = {"u": "u[{i}]", "p": "p[{i}]", "t": "t", "c": "c"}
    pr = Printer(g, fmt, interp_fmt="mb_interp_{kind}{deriv}(tb, {table}, {t})")
    args = "const double* MB_RESTRICT u, const double* MB_RESTRICT p, double t, const double* MB_RESTRICT tb"
    out: List[str] = []
    out.append(f"// GENERATED by multibody.jl_b200/codegen/model_emit.py -- do not edit.")
    out.append(f"// model '{X}': n={n} states, m={m} parameters, nnz(J)={code.nnz}, LU slots={code.nlu}")
    out.append(f"// algorithmic flops (BASELINE.md sec.4 table): rhs={code.F_rhs} jac={code.F_jac} "
               f"dT={code.F_dT} factor(jac+LU)={code.F_factor} solve={code.F_solve}")
    out.append("#pragma once")
    for nm in ("RHS", "FACTOR", "SOLVE"):      # per-function linkage override (tuning)
        out.append(f"#ifndef MB_FN_{nm}\n#define MB_FN_{nm} MB_FN\n#endif")
    out.append(f"#define MBM_{X}_N {n}")
    out.append(f"#define MBM_{X}_M {max(m, 0)}")
    out.append(f"#define MBM_{X}_NNZ {code.nnz}")
    out.append(f"#define MBM_{X}_NLU {code.nlu}")
    out.append(f"#define MBM_{X}_F_RHS {code.F_rhs}")
    out.append(f"#define MBM_{X}_F_JAC {code.F_jac}")
    out.append(f"#define MBM_{X}_F_DT {code.F_dT}")
    out.append(f"#define MBM_{X}_F_FACTOR {code.F_factor}")
    out.append(f"#define MBM_{X}_F_SOLVE {code.F_solve}")
    obs_names = sorted(sys.observed)
    out.append(f"#define MBM_{X}_NOBS {len(obs_names)}")
    slots = int(sys.meta.get("smem_slots", 0))
    out.append(f"#define MBM_{X}_SMEM_SLOTS {slots}")
    out.append(f"#define MBM_{X}_HAS_DT {1 if any(d != g.zero for d in code.dT_nodes) else 0}")
    # rows of df/dt that are not structurally zero (bit i; all ones when n > 64): the Rosenbrock row kernel keeps only those
    dt_mask = sum(1 << i for i, d in enumerate(code.dT_nodes) if d != g.zero) if n <= 64 else (1 << 64) - 1
    out.append(f"#define MBM_{X}_DT_MASK 0x{dt_mask:x}ull")
    out.append(f"#define MBM_{X}_IS_ODE {1 if all(sys.mass_diag) else 0}")
    out.append("#ifndef MB_NO_PATTERN")
    out.append(_carray("int", f"mbm_{X}_jrow", [e[0] for e in code.jac_entries]))
    out.append(_carray("int", f"mbm_{X}_jcol", [e[1] for e in code.jac_entries]))
    out.append(_carray("int", f"mbm_{X}_mass", [int(v) for v in sys.mass_diag]))
    out.append("#endif")
    words = []
    for w0 in range(0, n, 64):
        v = 0
        for i in range(w0, min(n, w0 + 64)):
            if sys.mass_diag[i]:
                v |= 1 << (i - w0)
        words.append(f"0x{v:x}ull")
    out.append(f"MB_FN int mbm_{X}_isdiff(int i) {{")
    out.append(f"  const unsigned long long w[{len(words)}] = {{{', '.join(words)}}};")
    out.append("  return (int)((w[i >> 6] >> (i & 63)) & 1ull);")
    out.append("}")
    out.append("")
    pr.smem_slots, pr.reg_target = slots, int(sys.meta.get("reg_target", 72))
    out.append(pr.function(f"MB_FN_RHS void mbm_{X}_rhs({args}, double* MB_RESTRICT du)",
                           [(f"du[{i}]", sys.rhs[i]) for i in range(n)]))
    if sys.meta.get("strided_rhs"):
        # second copy of the right-hand side reading its state from shared memory laid out
        # [component][thread] (u points at this thread's column): the stage state of the
        # Rosenbrock row kernel (csrc/mb_rodas_rows.cuh, MB_ROWS_SMEM = 2)
        pr.input_fmt = dict(fmt, u="u[{i} * MB_S_STRIDE]", p="MB_FP_AT({i})")
        out.append("#ifndef MB_FP_AT\n#define MB_FP_AT(i) p[i]   /* parameters as rhs_s / factor / dT / solve read them (a TU may keep them in shared memory) */\n#endif")
        out.append("#ifdef MB_RHS_STRIDED")
        out.append(pr.function(f"MB_FN_RHS void mbm_{X}_rhs_s(const double* u, const double* MB_RESTRICT p, double t, "
                               f"const double* MB_RESTRICT tb, double* MB_RESTRICT du)",
                               [(f"du[{i}]", sys.rhs[i]) for i in range(n)]))
        out.append("#endif")
        pr.input_fmt = fmt
    pr.smem_slots = 0
    out.append("")
    out.append(pr.function(f"MB_FN void mbm_{X}_jac({args}, double* MB_RESTRICT Jnz)",
                           [(f"Jnz[{k}]", e[2]) for k, e in enumerate(code.jac_entries)]))
    out.append("")
    # factor() and dT() read the state through MB_FU_AT so that a translation unit can keep it in shared memory, laid out
    # [component][thread] (Rosenbrock row kernel, MB_ROWS_SMEM = 6)
    out.append("#ifndef MB_FU_AT\n#define MB_FU_AT(i) u[i]\n#endif")
    out.append("#ifndef MB_FP_AT\n#define MB_FP_AT(i) p[i]\n#endif")
    pr.input_fmt = dict(fmt, u="MB_FU_AT({i})", p="MB_FP_AT({i})")
    out.append(pr.function(f"MB_FN void mbm_{X}_dT({args}, double* MB_RESTRICT dT)",
                           [(f"dT[{i}]", code.dT_nodes[i]) for i in range(n)]))
    pr.input_fmt = fmt
    out.append("")
    out.append(f"// observed variables, in this order: {', '.join(obs_names) if obs_names else '(none)'}")
    out.append(pr.function(f"MB_FN void mbm_{X}_observed({args}, double* MB_RESTRICT obs)",
                           [(f"obs[{k}]", sys.observed[nm]) for k, nm in enumerate(obs_names)]))
    out.append("")
    pr.smem_slots = slots if sys.meta.get("stage_factor", True) else 0
    out.append("#ifndef MB_LU_AT\n#define MB_LU_AT(k) LU[k]   /* slot k of this trajectory's LU storage (a TU may stride it) */\n#endif")
    lu_assign = [(f"MB_LU_AT({s})", node) for s, node in enumerate(code.lu_nodes)]
    pr.input_fmt = dict(fmt, u="MB_FU_AT({i})", p="MB_FP_AT({i})")
    out.append(pr.function(f"MB_FN_FACTOR void mbm_{X}_factor({args}, double c, double* MB_RESTRICT LU)",
                           lu_assign))
    pr.input_fmt = fmt
    out.append("")
    # runtime pivot guard: the pivot ORDER is fixed at code-generation time, so the code reports the largest stored
    # pivot reciprocal; the kernels retire a trajectory as Unstable when it is not finite or above MB_PIVOT_GUARD
    # (a numerically singular W, e.g. a loop in a singular configuration) instead of integrating on with a bad factorisation
    piv_slots = sorted({code.lu_slots[k] for k in code.lu_slots if k[0] == "d"})
    out.append(f"#define MBM_{X}_NPIV {len(piv_slots)}")
    pl = [f"MB_FN double mbm_{X}_pivmax(const double* MB_RESTRICT LU) {{", "  double g = 0.0;"]
    for sl in piv_slots:
        pl.append(f"  g = fmax(g, fabs(MB_LU_AT({sl})));")
    pl += ["  return g;", "}"]
    out.append("\n".join(pl))
    out.append("")
    pr.smem_slots = 0
    spr = Printer(code.solve_graph, {"lu": "MB_LU_AT({i})", "b": "b{i}", "c": "c", "p": "MB_FP_AT({i})"})
    sol_lines = [f"MB_FN_SOLVE void mbm_{X}_solve(const double* MB_RESTRICT LU, double c, const double* MB_RESTRICT p, "
                 f"double* MB_RESTRICT b) {{"]
    for i in range(n):
        sol_lines.append(f"  const double b{i} = b[{i}];")
    sol_lines += spr.body(code.solve_out)
    for i in range(n):
        sol_lines.append(f"  b[{i}] = {spr.ref(code.solve_out[i])};")
    sol_lines.append("}")
    out.append("\n".join(sol_lines))
    out.append("")
    return "\n".join(out)

"""DAG -> straight-line C / CUDA C printer.

The north star asks for a code generator that "takes the MTK-generated RHS,
mass matrix and Jacobian of the index-reduced system and emits straight-line
CUDA C compiled for sm_100a".  This module is that printer.  The same printer
emits plain C (used only by tests to check generated code on the CPU).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

from .dag import (Graph, CONST, INPUT, ADD, SUB, MUL, DIV, NEG, SIN, COS, SQRT,
                  ATAN2, ABS, MIN, MAX, SELGE, INTERP)


def _lit(v: float) -> str:
    if v == int(v) and abs(v) < 1e15:
        return f"{int(v)}.0"
    return repr(float(v))


class Printer:
    """Prints the nodes needed for a set of outputs as `const double xN = ...;`."""

    def __init__(self, g: Graph, input_fmt: Dict[str, str], interp_fmt: str =
                 "mb_interp_{kind}{deriv}(tb, {table}, {t})", scheduled: bool = True):
        self.g = g
        self.scheduled = scheduled
        self.input_fmt = input_fmt
        self.interp_fmt = interp_fmt
        # staged spilling: when > 0, `function` keeps at most `reg_target` temporaries in named
        # variables and parks the ones with the farthest next use in MB_S(slot) (shared memory
        # on the GPU), instead of leaving the choice -- and a trip to DRAM -- to ptxas
        self.smem_slots = 0
        self.reg_target = 72
        self._alias: Dict[int, str] = {}
        self._seg: Dict[int, str] = {}        # time node -> name of its table-segment variable
        self._paired: Dict[int, int] = {}     # sin/cos node -> partner emitted in the same MB_SINCOS
        self._emitted: set = set()

    def _begin(self, order: Sequence[int]) -> None:
        """Per-function state: table segments already located, sin/cos pairs of one argument."""
        g = self.g
        self._seg, self._paired, self._emitted = {}, {}, set()
        self._alias = {}
        by_arg: Dict[int, Dict[int, int]] = {}
        for n in order:
            if g.op[n] in (SIN, COS):
                by_arg.setdefault(g.a[n], {})[g.op[n]] = n
        for d in by_arg.values():
            if len(d) == 2:
                self._paired[d[SIN]] = d[COS]
                self._paired[d[COS]] = d[SIN]

    def ref(self, n: int) -> str:
        g = self.g
        o = g.op[n]
        if o == CONST:
            v = g.val[n]
            return _lit(v) if v >= 0 else f"({_lit(v)})"
        if o == INPUT:
            kind, idx = g.payload[n]
            return self.input_fmt[kind].format(i=idx)
        if o == NEG:
            return f"(-{self.ref(g.a[n])})"
        return self._alias.get(n) or f"x{n}"

    def _line(self, n: int, indent: str):
        g = self.g
        r = self.ref
        o = g.op[n]
        if o in (CONST, INPUT, NEG):
            return None
        if n in self._emitted:            # second half of a sin/cos pair
            return None
        a, b, c = g.a[n], g.b[n], g.c[n]
        if o in (SIN, COS) and n in self._paired:
            m = self._paired[n]
            self._emitted.add(m)
            sn, cn = (n, m) if o == SIN else (m, n)
            return f"{indent}double x{sn}, x{cn}; MB_SINCOS({r(a)}, x{sn}, x{cn});"
        if o == INTERP:
            kind, table, deriv = g.payload[n]
            pre = ""
            if a not in self._seg:
                self._seg[a] = f"sg{a}"
                pre = f"{indent}const MB_SEG_T sg{a} = MB_SEG(tb, {r(a)});\n"
            return (f"{pre}{indent}const double x{n} = MB_TAB_{kind.upper()}{deriv}(tb, {self._seg[a]}, {table});")
        if o in (ADD, SUB, MUL, DIV):
            e = f"{r(a)} {'+-*/'[o - ADD]} {r(b)}"
        elif o == SIN:
            e = f"sin({r(a)})"
        elif o == COS:
            e = f"cos({r(a)})"
        elif o == SQRT:
            e = f"sqrt({r(a)})"
        elif o == ABS:
            e = f"fabs({r(a)})"
        elif o == ATAN2:
            e = f"atan2({r(a)}, {r(b)})"
        elif o == MIN:
            e = f"fmin({r(a)}, {r(b)})"
        elif o == MAX:
            e = f"fmax({r(a)}, {r(b)})"
        elif o == SELGE:
            e = f"(({r(a)} >= 0.0) ? {r(b)} : {r(c)})"
        elif o == INTERP:
            kind, table, deriv = g.payload[n]
            e = self.interp_fmt.format(kind=kind, deriv=deriv, table=table, t=r(a))
        else:
            raise NotImplementedError(o)
        return f"{indent}const double x{n} = {e};"

    def body(self, outs: Sequence[int], indent: str = "  ") -> List[str]:
        g = self.g
        lines: List[str] = []
        order = g.schedule(outs) if self.scheduled else g.reachable(outs)
        self._begin(order)
        for n in order:
            ln = self._line(n, indent)
            if ln is not None:
                lines.append(ln)
        return lines

    def function(self, signature: str, assigns: Sequence[Tuple[str, int]],
                 indent: str = "  ") -> str:
        """`assigns` = [(lhs C expression, node)].  With scheduling on, each output is
        stored as soon as its value exists (short live ranges); otherwise all stores
        follow the temporaries."""
        g = self.g
        lines = [signature + " {"]
        self._begin(g.reachable([n for _, n in assigns]))
        if not self.scheduled:
            lines += self.body([n for _, n in assigns], indent)
            for lhs, n in assigns:
                lines.append(f"{indent}{lhs} = {self.ref(n)};")
        else:
            order = g.schedule([n for _, n in assigns])
            by_node: Dict[int, List[str]] = {}
            for lhs, n in assigns:
                by_node.setdefault(n, []).append(lhs)
            emitted_lhs = set()
            stage = self.smem_slots > 0
            if stage:
                import bisect
                trivial = (CONST, INPUT, NEG)

                def operands(n):
                    o = g.op[n]
                    if o in (CONST, INPUT):
                        return []
                    ks = [g.a[n]]
                    if o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
                        ks.append(g.b[n])
                    elif o == SELGE:
                        ks += [g.b[n], g.c[n]]
                    out = []
                    for k in ks:
                        while g.op[k] == NEG:
                            k = g.a[k]
                        if g.op[k] not in (CONST, INPUT):
                            out.append(k)
                    return out
                uses: Dict[int, List[int]] = {}
                for i, n in enumerate(order):
                    if g.op[n] in trivial:
                        continue
                    for k in operands(n):
                        uses.setdefault(k, []).append(i)

                def next_use(k, i):
                    u = uses.get(k, ())
                    j = bisect.bisect_right(u, i)
                    return u[j] if j < len(u) else None
                live: set = set()
                slot_of: Dict[int, int] = {}
                free = list(range(self.smem_slots - 1, -1, -1))
                nreload = 0
                lines.append(f"{indent}MB_S_DECL")
            for i, n in enumerate(order):
                if stage and g.op[n] not in (CONST, INPUT, NEG) and n not in self._emitted:
                    for k in operands(n):                     # bring parked operands back
                        if k not in live and k in slot_of:
                            nreload += 1
                            nm = f"x{k}_r{nreload}"
                            lines.append(f"{indent}const double {nm} = MB_S({slot_of[k]});")
                            self._alias[k] = nm
                            live.add(k)
                ln = self._line(n, indent)
                if ln is not None:
                    lines.append(ln)
                for lhs in by_node.get(n, ()):
                    lines.append(f"{indent}{lhs} = {self.ref(n)};")
                    emitted_lhs.add(lhs)
                if stage and g.op[n] not in (CONST, INPUT, NEG):
                    if ln is not None:
                        for d_ in (n, self._paired.get(n)):
                            if d_ is not None and next_use(d_, i) is not None:
                                live.add(d_)
                    for k in list(live):                      # values that are dead from here on
                        if next_use(k, i) is None:
                            live.discard(k)
                            if k in slot_of:
                                free.append(slot_of.pop(k))
                    nxt = set(operands(order[i + 1])) if i + 1 < len(order) else set()
                    while len(live) > self.reg_target:
                        cand = [k for k in live if k not in nxt]
                        if not cand:
                            break
                        v = max(cand, key=lambda k: next_use(k, i))
                        if v not in slot_of:
                            if not free:
                                break
                            slot_of[v] = free.pop()
                            lines.append(f"{indent}MB_S({slot_of[v]}) = {self.ref(v)};")
                        live.discard(v)
            for lhs, n in assigns:          # outputs that are constants / inputs
                if lhs not in emitted_lhs:
                    lines.append(f"{indent}{lhs} = {self.ref(n)};")
        lines.append("}")
        return "\n".join(lines)


# ---------------------------------------------------------------------------
# codegen-time sparse LU (pivot sequence fixed at code-generation time)
# ---------------------------------------------------------------------------

def sparse_lu(g: Graph, W: Dict[Tuple[int, int], int], n: int, pivots=None):
    """Symbolic LU of the sparse matrix W (dict (i,j)->node).

    `pivots` is a sequence of (row, col) chosen at code-generation time (see
    `choose_pivots_numeric`); when None, diagonal pivots in Markowitz (min
    fill) order are used.

    Returns (order, invd, L, U), all indexed by elimination step s:
    order[s] = (prow, pcol); invd[s] node of 1/pivot; L[s] = [(i, node)]
    multipliers of the rows eliminated with that pivot; U[s] = [(j, node)]
    remaining entries of the pivot row.  W is consumed.
    """
    rows: Dict[int, Dict[int, int]] = {i: {} for i in range(n)}
    cols: Dict[int, set] = {j: set() for j in range(n)}
    for (i, j), v in W.items():
        if v == g.zero:
            continue
        rows[i][j] = v
        cols[j].add(i)
    rrem = set(range(n))
    crem = set(range(n))
    order: List[Tuple[int, int]] = []
    invd: Dict[int, int] = {}
    L: Dict[int, List[Tuple[int, int]]] = {}
    U: Dict[int, List[Tuple[int, int]]] = {}
    for step in range(n):
        if pivots is not None:
            pr, pc = pivots[step]
            if pc not in rows[pr]:
                raise ValueError(f"pivot ({pr},{pc}) is structurally zero")
        else:
            best, bestcost = None, None
            for k in sorted(rrem):
                if k not in rows[k] or k not in crem:
                    continue
                rc = sum(1 for j in rows[k] if j in crem) - 1
                cc = sum(1 for i in cols[k] if i in rrem) - 1
                cost = rc * cc
                if bestcost is None or cost < bestcost:
                    best, bestcost = k, cost
            if best is None:
                raise ValueError("structurally zero diagonal: static diagonal pivoting impossible")
            pr = pc = best
        rrem.discard(pr)
        crem.discard(pc)
        order.append((pr, pc))
        inv = g.div(g.one, rows[pr][pc])
        invd[step] = inv
        urow = [(j, v) for j, v in sorted(rows[pr].items()) if j in crem]
        U[step] = urow
        lk: List[Tuple[int, int]] = []
        for i in sorted(cols[pc]):
            if i not in rrem or pc not in rows[i]:
                continue
            l = g.mul(rows[i][pc], inv)
            if l == g.zero:
                continue
            lk.append((i, l))
            ri = rows[i]
            for j, v in urow:
                upd = g.mul(l, v)
                if j in ri:
                    ri[j] = g.sub(ri[j], upd)
                else:
                    ri[j] = g.neg(upd)
                    cols[j].add(i)
        L[step] = lk
    return order, invd, L, U


def lu_solve(g: Graph, order, invd, L, U, b: Sequence[int]) -> List[int]:
    """x = W^-1 b using factors from `sparse_lu` (all node ids)."""
    y = list(b)
    for s, (pr, pc) in enumerate(order):
        yk = y[pr]
        if yk == g.zero:
            continue
        for i, l in L[s]:
            y[i] = g.sub(y[i], g.mul(l, yk))
    x = [g.zero] * len(order)
    for s in reversed(range(len(order))):
        pr, pc = order[s]
        acc = y[pr]
        for j, v in U[s]:
            acc = g.sub(acc, g.mul(v, x[j]))
        x[pc] = g.mul(acc, invd[s])
    return x


def choose_pivots_numeric(Wnum, thresh: float = 0.1):
    """Pivot sequence for `sparse_lu` from a numeric sample of W (dense ndarray,
    structural zeros exactly 0): Markowitz cost with threshold partial pivoting
    (|a_ij| >= thresh * max|a_.j| in its column), fill-in tracked on the sample."""
    import numpy as np
    A = np.array(Wnum, dtype=float)
    n = A.shape[0]
    nzmask = A != 0.0
    rrem = list(range(n)); crem = list(range(n))
    seq = []
    for _ in range(n):
        best = None
        for j in crem:
            colmax = max(abs(A[i, j]) for i in rrem)
            if colmax == 0.0:
                continue
            cc = sum(1 for i in rrem if nzmask[i, j]) - 1
            for i in rrem:
                if not nzmask[i, j] or abs(A[i, j]) < thresh * colmax:
                    continue
                rc = sum(1 for jj in crem if nzmask[i, jj]) - 1
                key = (rc * cc, -abs(A[i, j]) / colmax, i != j, i, j)
                if best is None or key < best[0]:
                    best = (key, i, j)
        if best is None:
            raise ValueError("numerically singular sample matrix")
        _, pr, pc = best
        seq.append((pr, pc))
        rrem.remove(pr); crem.remove(pc)
        for i in rrem:
            if nzmask[i, pc]:
                l = A[i, pc] / A[pr, pc]
                for j in crem:
                    if nzmask[pr, j]:
                        A[i, j] -= l * A[pr, j]
                        nzmask[i, j] = True
    return seq


def ldl_solve_factory(g: Graph, M: List[List[int]]):
    """Straight-line LDL^T of a small dense SPD matrix (node ids); returns a
    function rhs(list of nodes) -> solution nodes.  Used for the linear
    subsystems MTK inlines into the RHS (reference: src/Multibody.jl:159-166,
    `inline_linear_sccs=true`), e.g. the joint-space mass matrix."""
    n = len(M)
    Lm = [[g.zero] * n for _ in range(n)]
    Dinv = [g.zero] * n
    Dd = [g.zero] * n
    for j in range(n):
        d = M[j][j]
        for k in range(j):
            d = g.sub(d, g.mul(g.mul(Lm[j][k], Lm[j][k]), Dd[k]))
        Dd[j] = d
        Dinv[j] = g.div(g.one, d)
        for i in range(j + 1, n):
            s = M[i][j]
            for k in range(j):
                s = g.sub(s, g.mul(g.mul(Lm[i][k], Lm[j][k]), Dd[k]))
            Lm[i][j] = g.mul(s, Dinv[j])

    def solve(b: Sequence[int]) -> List[int]:
        y = list(b)
        for i in range(n):
            for k in range(i):
                y[i] = g.sub(y[i], g.mul(Lm[i][k], y[k]))
        z = [g.mul(y[i], Dinv[i]) for i in range(n)]
        x = list(z)
        for i in reversed(range(n)):
            for k in range(i + 1, n):
                x[i] = g.sub(x[i], g.mul(Lm[k][i], x[k]))
        return x

    return solve

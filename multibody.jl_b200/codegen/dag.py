"""Hash-consed scalar expression DAG over f64: the IR of the B200 codegen backend.

This is the stand-in for the symbolic RHS / mass matrix / Jacobian that
ModelingToolkit's `mtkcompile` hands to its code generator
(reference call site: src/Multibody.jl:184).  A model frontend builds nodes
through ordinary Python arithmetic; the backend differentiates them
(forward mode, sparsity aware), counts their algorithmic flops with the cost
table fixed in BASELINE.md §4, and prints straight-line C / CUDA C.

Nodes are integers into flat arrays (fast, no per-node objects on hot paths);
`Ex` is a thin operator-overloading handle around a node id.
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List, Sequence, Tuple

# op codes
CONST, INPUT, ADD, SUB, MUL, DIV, NEG, SIN, COS, SQRT, ATAN2, ABS, MIN, MAX, \
    SELGE, INTERP = range(16)

_OPNAME = {ADD: "+", SUB: "-", MUL: "*", DIV: "/"}

# BASELINE.md §4 cost table (flops per op)
FLOP_COST = {CONST: 0, INPUT: 0, ADD: 1, SUB: 1, MUL: 1, DIV: 4, NEG: 0, SIN: 20,
             COS: 20, SQRT: 4, ATAN2: 30, ABS: 1, MIN: 1, MAX: 1, SELGE: 1,
             INTERP: 12}


class Graph:
    """One expression graph (one model).  All `Ex` of a model share it."""

    def __init__(self):
        self.op: List[int] = []
        self.a: List[int] = []
        self.b: List[int] = []
        self.c: List[int] = []          # third operand (SELGE) / payload index
        self.val: List[float] = []      # CONST value
        self.payload: List[tuple] = []  # INPUT: (kind, idx); INTERP: (kind, table, deriv)
        self._memo: Dict[tuple, int] = {}
        self.zero = self.const(0.0)
        self.one = self.const(1.0)

    # ---- node construction -------------------------------------------------
    def _new(self, op, a=-1, b=-1, c=-1, val=0.0, payload=None, key=None):
        key = key if key is not None else (op, a, b, c)
        n = self._memo.get(key)
        if n is not None:
            return n
        n = len(self.op)
        self.op.append(op); self.a.append(a); self.b.append(b); self.c.append(c)
        self.val.append(val); self.payload.append(payload)
        self._memo[key] = n
        return n

    def const(self, v: float) -> int:
        v = float(v)
        if v == 0.0:
            v = 0.0  # fold -0.0
        return self._new(CONST, val=v, key=(CONST, v))

    def input(self, kind: str, idx: int) -> int:
        return self._new(INPUT, payload=(kind, idx), key=(INPUT, kind, idx))

    def isconst(self, n: int) -> bool:
        return self.op[n] == CONST

    def add(self, x: int, y: int) -> int:
        op = self.op
        if op[x] == CONST and op[y] == CONST:
            return self.const(self.val[x] + self.val[y])
        if op[x] == CONST and self.val[x] == 0.0:
            return y
        if op[y] == CONST and self.val[y] == 0.0:
            return x
        if op[y] == NEG:
            return self.sub(x, self.a[y])
        if op[x] == NEG:
            return self.sub(y, self.a[x])
        if x > y:
            x, y = y, x
        return self._new(ADD, x, y)

    def sub(self, x: int, y: int) -> int:
        op = self.op
        if op[x] == CONST and op[y] == CONST:
            return self.const(self.val[x] - self.val[y])
        if op[y] == CONST and self.val[y] == 0.0:
            return x
        if op[x] == CONST and self.val[x] == 0.0:
            return self.neg(y)
        if x == y:
            return self.zero
        if op[y] == NEG:
            return self.add(x, self.a[y])
        return self._new(SUB, x, y)

    def neg(self, x: int) -> int:
        op = self.op
        if op[x] == CONST:
            return self.const(-self.val[x])
        if op[x] == NEG:
            return self.a[x]
        if op[x] == SUB:
            return self.sub(self.b[x], self.a[x])
        return self._new(NEG, x)

    def mul(self, x: int, y: int) -> int:
        op = self.op
        if op[x] == CONST and op[y] == CONST:
            return self.const(self.val[x] * self.val[y])
        if op[y] == CONST:
            x, y = y, x
        if op[x] == CONST:
            v = self.val[x]
            if v == 0.0:
                return self.zero
            if v == 1.0:
                return y
            if v == -1.0:
                return self.neg(y)
            if op[y] == NEG:
                return self.mul(self.const(-v), self.a[y])
            if op[y] == MUL and op[self.a[y]] == CONST:
                return self.mul(self.const(v * self.val[self.a[y]]), self.b[y])
            return self._new(MUL, x, y)  # const always first
        if op[x] == NEG and op[y] == NEG:
            return self.mul(self.a[x], self.a[y])
        if op[x] == NEG:
            return self.neg(self.mul(self.a[x], y))
        if op[y] == NEG:
            return self.neg(self.mul(x, self.a[y]))
        if x > y:
            x, y = y, x
        return self._new(MUL, x, y)

    def div(self, x: int, y: int) -> int:
        op = self.op
        if op[y] == CONST:
            if self.val[y] == 0.0:
                raise ZeroDivisionError("symbolic division by constant zero")
            return self.mul(self.const(1.0 / self.val[y]), x)
        if op[x] == CONST and self.val[x] == 0.0:
            return self.zero
        if op[x] == NEG:
            return self.neg(self.div(self.a[x], y))
        if op[y] == NEG:
            return self.neg(self.div(x, self.a[y]))
        return self._new(DIV, x, y)

    def unary(self, opc: int, x: int) -> int:
        if self.op[x] == CONST:
            v = self.val[x]
            f = {SIN: math.sin, COS: math.cos, SQRT: math.sqrt, ABS: abs}[opc]
            return self.const(f(v))
        if opc == SIN and self.op[x] == NEG:
            return self.neg(self.unary(SIN, self.a[x]))
        if opc in (COS, ABS) and self.op[x] == NEG:
            return self.unary(opc, self.a[x])
        return self._new(opc, x)

    def binary(self, opc: int, x: int, y: int) -> int:
        if self.op[x] == CONST and self.op[y] == CONST:
            f = {ATAN2: math.atan2, MIN: min, MAX: max}[opc]
            return self.const(f(self.val[x], self.val[y]))
        if opc in (MIN, MAX) and x > y:
            x, y = y, x
        return self._new(opc, x, y)

    def selge(self, cond: int, x: int, y: int) -> int:
        """cond >= 0 ? x : y"""
        if self.op[cond] == CONST:
            return x if self.val[cond] >= 0.0 else y
        if x == y:
            return x
        return self._new(SELGE, cond, x, y)

    def interp(self, kind: str, table: int, deriv: int, t: int) -> int:
        """Shared time table lookup (reference: DataInterpolations objects baked
        into parameters at src/robot/path_planning.jl:191-193).
        kind: 'cubic' | 'linear'; deriv: 0 value, 1 d/dt, 2 d2/dt2."""
        return self._new(INTERP, t, payload=(kind, table, deriv),
                         key=(INTERP, t, kind, table, deriv))

    # ---- analysis ----------------------------------------------------------
    def reachable(self, outs: Iterable[int]) -> List[int]:
        """Node ids needed for `outs`, ascending (= topological) order."""
        seen = set()
        stack = [o for o in outs]
        op, a, b, c = self.op, self.a, self.b, self.c
        while stack:
            n = stack.pop()
            if n in seen:
                continue
            seen.add(n)
            o = op[n]
            if o in (CONST, INPUT):
                continue
            stack.append(a[n])
            if o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
                stack.append(b[n])
            elif o == SELGE:
                stack.append(b[n]); stack.append(c[n])
        return sorted(seen)

    def schedule(self, outs: Sequence[int]) -> List[int]:
        """Evaluation order for straight-line code: depth-first post-order from the
        outputs, heavier operand first (Sethi-Ullman flavoured).  Compared with plain
        creation order this keeps producers next to their consumers, which shortens
        live ranges and cuts register spills in the emitted CUDA."""
        op, a, b, c = self.op, self.a, self.b, self.c
        need = self.reachable(outs)
        weight: Dict[int, int] = {}
        for n in need:                       # ascending = topological
            o = op[n]
            if o in (CONST, INPUT):
                weight[n] = 0
            elif o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
                weight[n] = 1 + weight[a[n]] + weight[b[n]]
            elif o == SELGE:
                weight[n] = 1 + weight[a[n]] + weight[b[n]] + weight[c[n]]
            else:
                weight[n] = 1 + weight[a[n]]
            if weight[n] > 1 << 40:
                weight[n] = 1 << 40
        done = set()
        order: List[int] = []
        for root in outs:
            stack = [(root, False)]
            while stack:
                n, expanded = stack.pop()
                if n in done:
                    continue
                o = op[n]
                if o in (CONST, INPUT):
                    done.add(n); order.append(n)
                    continue
                if expanded:
                    done.add(n); order.append(n)
                    continue
                stack.append((n, True))
                if o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
                    kids = [a[n], b[n]]
                elif o == SELGE:
                    kids = [a[n], b[n], c[n]]
                else:
                    kids = [a[n]]
                kids.sort(key=lambda k: weight[k])      # lightest pushed first -> heaviest visited first
                for kid in kids:
                    if kid not in done:
                        stack.append((kid, False))
        return order

    def flops(self, outs: Iterable[int]) -> int:
        """Algorithmic flop count.  An ADD/SUB with a MUL
        operand still counts 2 in total (fma = 2), so no special case."""
        return sum(FLOP_COST[self.op[n]] for n in self.reachable(outs))

    def nnodes(self, outs: Iterable[int]) -> int:
        return sum(1 for n in self.reachable(outs) if self.op[n] not in (CONST, INPUT))

    # ---- forward-mode differentiation (sparse) ----------------------------
    def diff(self, outs: Sequence[int], wrt: Sequence[int]) -> List[Dict[int, int]]:
        """d outs[i] / d wrt[j] as a list (per out) of {j: node}.  `wrt` are
        INPUT node ids.  Only structurally non-zero entries are stored."""
        order = self.reachable(outs)
        wpos = {w: j for j, w in enumerate(wrt)}
        op, a, b, c = self.op, self.a, self.b, self.c
        D: Dict[int, Dict[int, int]] = {}
        empty: Dict[int, int] = {}

        def lin(da, ca, db, cb):
            """ca*da + cb*db over sparse dicts; ca/cb node ids or None (=1) / 'neg'."""
            out: Dict[int, int] = {}
            for j, d in da.items():
                out[j] = d if ca is None else self.mul(ca, d)
            for j, d in db.items():
                t = d if cb is None else self.mul(cb, d)
                out[j] = self.add(out[j], t) if j in out else t
            return {j: d for j, d in out.items() if d != self.zero}

        for n in order:
            o = op[n]
            if o == CONST:
                continue
            if o == INPUT:
                if n in wpos:
                    D[n] = {wpos[n]: self.one}
                continue
            da = D.get(a[n], empty)
            if o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
                db = D.get(b[n], empty)
                if not da and not db:
                    continue
                x, y = a[n], b[n]
                if o == ADD:
                    r = lin(da, None, db, None)
                elif o == SUB:
                    r = lin(da, None, db, self.const(-1.0))
                elif o == MUL:
                    r = lin(da, y, db, x)
                elif o == DIV:
                    inv = self.div(self.one, y)
                    r = lin(da, inv, db, self.neg(self.mul(n, inv)))
                elif o == ATAN2:  # atan2(x=y-arg first, y)
                    den = self.add(self.mul(x, x), self.mul(y, y))
                    r = lin(da, self.div(y, den), db, self.neg(self.div(x, den)))
                elif o == MIN:
                    cnd = self.sub(y, x)  # >=0 -> x chosen
                    r = self._sel_d(cnd, da, db)
                else:  # MAX
                    cnd = self.sub(x, y)
                    r = self._sel_d(cnd, da, db)
            elif o == SELGE:
                db = D.get(b[n], empty); dc = D.get(c[n], empty)
                if not db and not dc:
                    continue
                r = self._sel_d(a[n], db, dc)
            else:
                if not da:
                    continue
                x = a[n]
                if o == NEG:
                    r = {j: self.neg(d) for j, d in da.items()}
                elif o == SIN:
                    r = lin(da, self.unary(COS, x), empty, None)
                elif o == COS:
                    r = lin(da, self.neg(self.unary(SIN, x)), empty, None)
                elif o == SQRT:
                    r = lin(da, self.div(self.const(0.5), n), empty, None)
                elif o == ABS:
                    sgn = self.selge(x, self.one, self.const(-1.0))
                    r = lin(da, sgn, empty, None)
                elif o == INTERP:
                    kind, table, deriv = self.payload[n]
                    r = lin(da, self.interp(kind, table, deriv + 1, x), empty, None)
                else:
                    raise NotImplementedError(o)
            if r:
                D[n] = r
        return [D.get(o, {}) if self.op[o] != INPUT or o not in wpos
                else {wpos[o]: self.one} for o in outs]

    def _sel_d(self, cnd, dx, dy):
        out = {}
        for j in set(dx) | set(dy):
            out[j] = self.selge(cnd, dx.get(j, self.zero), dy.get(j, self.zero))
        return {j: d for j, d in out.items() if d != self.zero}

    # ---- substitution -----------------------------------------------------
    def subs(self, outs: Sequence[int], mapping: Dict[int, int]) -> List[int]:
        """Rebuild `outs` with INPUT nodes replaced per `mapping` (node -> node)."""
        order = self.reachable(outs)
        new: Dict[int, int] = {}
        op, a, b, c = self.op, self.a, self.b, self.c
        for n in order:
            o = op[n]
            if n in mapping:
                new[n] = mapping[n]
            elif o in (CONST, INPUT):
                new[n] = n
            elif o == ADD:
                new[n] = self.add(new[a[n]], new[b[n]])
            elif o == SUB:
                new[n] = self.sub(new[a[n]], new[b[n]])
            elif o == MUL:
                new[n] = self.mul(new[a[n]], new[b[n]])
            elif o == DIV:
                new[n] = self.div(new[a[n]], new[b[n]])
            elif o == NEG:
                new[n] = self.neg(new[a[n]])
            elif o in (SIN, COS, SQRT, ABS):
                new[n] = self.unary(o, new[a[n]])
            elif o in (ATAN2, MIN, MAX):
                new[n] = self.binary(o, new[a[n]], new[b[n]])
            elif o == SELGE:
                new[n] = self.selge(new[a[n]], new[b[n]], new[c[n]])
            elif o == INTERP:
                k, tb, dv = self.payload[n]
                new[n] = self.interp(k, tb, dv, new[a[n]])
        return [new[o] for o in outs]

    # ---- numeric evaluation (tests, CPU-side checks of generated code) ------
    def evaluate(self, outs: Sequence[int], inputs: Dict[Tuple[str, int], float],
                 tables=None) -> List[float]:
        order = self.reachable(outs)
        v: Dict[int, float] = {}
        op, a, b, c = self.op, self.a, self.b, self.c
        for n in order:
            o = op[n]
            if o == CONST:
                v[n] = self.val[n]
            elif o == INPUT:
                v[n] = float(inputs[self.payload[n]])
            elif o == ADD:
                v[n] = v[a[n]] + v[b[n]]
            elif o == SUB:
                v[n] = v[a[n]] - v[b[n]]
            elif o == MUL:
                v[n] = v[a[n]] * v[b[n]]
            elif o == DIV:
                v[n] = v[a[n]] / v[b[n]]
            elif o == NEG:
                v[n] = -v[a[n]]
            elif o == SIN:
                v[n] = math.sin(v[a[n]])
            elif o == COS:
                v[n] = math.cos(v[a[n]])
            elif o == SQRT:
                v[n] = math.sqrt(v[a[n]])
            elif o == ABS:
                v[n] = abs(v[a[n]])
            elif o == ATAN2:
                v[n] = math.atan2(v[a[n]], v[b[n]])
            elif o == MIN:
                v[n] = min(v[a[n]], v[b[n]])
            elif o == MAX:
                v[n] = max(v[a[n]], v[b[n]])
            elif o == SELGE:
                v[n] = v[b[n]] if v[a[n]] >= 0.0 else v[c[n]]
            elif o == INTERP:
                k, tb, dv = self.payload[n]
                v[n] = tables.eval(k, tb, dv, v[a[n]])
        return [v[o] for o in outs]


class Ex:
    """Operator-overloading handle: `Ex(g, node)`."""
    __slots__ = ("g", "n")
    __array_priority__ = 1000  # make numpy defer to our reflected ops

    def __init__(self, g: Graph, n: int):
        self.g = g
        self.n = n

    def _w(self, o) -> int:
        if isinstance(o, Ex):
            return o.n
        return self.g.const(float(o))

    def __add__(self, o): return Ex(self.g, self.g.add(self.n, self._w(o)))
    def __radd__(self, o): return Ex(self.g, self.g.add(self._w(o), self.n))
    def __sub__(self, o): return Ex(self.g, self.g.sub(self.n, self._w(o)))
    def __rsub__(self, o): return Ex(self.g, self.g.sub(self._w(o), self.n))
    def __mul__(self, o): return Ex(self.g, self.g.mul(self.n, self._w(o)))
    def __rmul__(self, o): return Ex(self.g, self.g.mul(self._w(o), self.n))
    def __truediv__(self, o): return Ex(self.g, self.g.div(self.n, self._w(o)))
    def __rtruediv__(self, o): return Ex(self.g, self.g.div(self._w(o), self.n))
    def __neg__(self): return Ex(self.g, self.g.neg(self.n))
    def __pos__(self): return self

    def is_zero(self) -> bool:
        return self.n == self.g.zero

    def __repr__(self):
        g = self.g
        if g.op[self.n] == CONST:
            return f"Ex({g.val[self.n]!r})"
        return f"Ex(#{self.n})"


def sin(x: Ex) -> Ex: return Ex(x.g, x.g.unary(SIN, x.n))
def cos(x: Ex) -> Ex: return Ex(x.g, x.g.unary(COS, x.n))
def sqrt(x: Ex) -> Ex: return Ex(x.g, x.g.unary(SQRT, x.n))
def fabs(x: Ex) -> Ex: return Ex(x.g, x.g.unary(ABS, x.n))
def atan2(y: Ex, x: Ex) -> Ex: return Ex(y.g, y.g.binary(ATAN2, y.n, x.n))


def fmax(x: Ex, y) -> Ex:
    return Ex(x.g, x.g.binary(MAX, x.n, x._w(y)))


def fmin(x: Ex, y) -> Ex:
    return Ex(x.g, x.g.binary(MIN, x.n, x._w(y)))


def select_ge(cond: Ex, x, y) -> Ex:
    return Ex(cond.g, cond.g.selge(cond.n, cond._w(x), cond._w(y)))

"""Warp-team code generation: one trajectory evaluated by G cooperating warps.

North star (2): "one trajectory per thread (or per warp for large chains/ropes) ... the small
dense LU factorisation done in registers or shared memory".  For n >= 30 a single thread cannot
keep the Rosenbrock work vectors on chip (robot: 7.9 KB of thread-local state per trajectory,
which round 1 measured as 56 MB of DRAM spill traffic per trajectory).  Here the straight-line
DAG of a generated function (rhs, W-build + LU `factor`, `solve`, df/dt) is split over G *roles*;
role w is executed by warp w of a CTA, lane l of every warp works on trajectory l of the CTA's
32 trajectories.  All lanes of a warp run the same instruction (no SIMT divergence: different
roles are different warps), and the per-trajectory state is spread over G register files plus
shared memory.

The split is a bulk-synchronous schedule: supersteps separated by a CTA barrier; inside a
superstep a role only uses values it computed itself or values other roles published to the
shared exchange area `X` in EARLIER supersteps.  `partition` grows, per superstep, G disjoint
ancestor-closed cones of the remaining DAG (list scheduling by critical path, least-loaded role
first), which keeps chains of dependent operations on one role and cuts only where the DAG
fans out.  `simulate` replays a plan numerically (role-private values + exchange slots with
their reuse) against the plain evaluation of the DAG; tests/test_team.py runs it for every
function of the robot.
"""
from __future__ import annotations

import heapq
from typing import Dict, List, Optional, Sequence, Tuple

from .dag import (Graph, CONST, INPUT, ADD, SUB, MUL, DIV, NEG, SIN, COS, SQRT, ATAN2, ABS, MIN,
                  MAX, SELGE, INTERP)
from .emit import Printer

# issue-slot cost model of the partitioner (FP64 pipe: one warp instruction per 2 cycles; the
# special functions are short polynomial sequences)
COST = {CONST: 0, INPUT: 0, ADD: 1, SUB: 1, MUL: 1, DIV: 8, NEG: 0, SIN: 25, COS: 25, SQRT: 8, ATAN2: 40,
        ABS: 1, MIN: 1, MAX: 1, SELGE: 1, INTERP: 8}
BARRIER_COST = 20      # cycles a superstep boundary costs every role
SEND_COST = 1          # one shared-memory store per published value
RECV_COST = 1          # one shared-memory load per consumed value


def _kids(g: Graph, n: int) -> List[int]:
    o = g.op[n]
    if o in (CONST, INPUT):
        return []
    ks = [g.a[n]]
    if o in (ADD, SUB, MUL, DIV, ATAN2, MIN, MAX):
        ks.append(g.b[n])
    elif o == SELGE:
        ks += [g.b[n], g.c[n]]
    return ks


def _base(g: Graph, k: int) -> int:
    while g.op[k] == NEG:
        k = g.a[k]
    return k


class TeamPlan:
    """Result of `partition`: who computes what, when, and what crosses roles."""

    def __init__(self, g: Graph, outs: Sequence[int], G: int):
        self.g, self.outs, self.G = g, list(outs), G
        self.proc: Dict[int, int] = {}
        self.step: Dict[int, int] = {}
        self.nsteps = 0
        self.blocks: List[List[List[int]]] = []          # [step][role] -> nodes in emission order
        self.slot: Dict[int, int] = {}                   # published node -> exchange slot
        self.recv: List[List[List[int]]] = []            # [step][role] -> nodes to read from X at block start
        self.send: List[List[List[int]]] = []            # [step][role] -> nodes to publish at block end
        self.out_role: List[int] = []                    # role that stores output i
        self.out_step: List[int] = []
        self.nslots = 0
        self.loads: List[List[int]] = []

    # ---- figures of merit -------------------------------------------------------------------
    @property
    def work(self) -> int:
        return sum(COST[self.g.op[n]] for n in self.proc)

    @property
    def makespan(self) -> int:
        return sum(max(l) for l in self.loads) + BARRIER_COST * self.nsteps

    @property
    def speedup(self) -> float:
        return self.work / max(self.makespan, 1)


def partition(g: Graph, outs: Sequence[int], G: int, thickness: int = 250,
              pin: Optional[Dict[int, int]] = None) -> TeamPlan:
    """Bulk-synchronous split of the DAG of `outs` over G roles.  `thickness` caps the work a
    role takes in one superstep (smaller: better balance, more barriers).  `pin` forces nodes
    onto a role."""
    plan = TeamPlan(g, outs, G)
    pin = pin or {}
    order = g.reachable(outs)
    work = [n for n in order if g.op[n] not in (CONST, INPUT, NEG)]
    wset = set(work)
    preds = {n: sorted(set(_base(g, k) for k in _kids(g, n)) & wset) for n in work}
    # sin and cos of one argument are emitted as one sincos: keep them on one role by making
    # the later node depend on the earlier one
    by_arg: Dict[int, Dict[int, int]] = {}
    for n in work:
        if g.op[n] in (SIN, COS):
            by_arg.setdefault(g.a[n], {})[g.op[n]] = n
    for d in by_arg.values():
        if len(d) == 2:
            lo, hi = sorted(d.values())
            if lo not in preds[hi]:
                preds[hi] = sorted(preds[hi] + [lo])
    succ: Dict[int, List[int]] = {n: [] for n in work}
    for n in work:
        for k in preds[n]:
            succ[k].append(n)
    bl: Dict[int, int] = {}                         # bottom level = critical path to a sink
    for n in reversed(work):
        bl[n] = COST[g.op[n]] + max((bl[m] for m in succ[n]), default=0)
    unsat = {n: [len(preds[n])] * G for n in work}
    remaining = set(work)
    proc, step = plan.proc, plan.step
    s = 0
    while remaining:
        load = [0] * G
        heaps: List[list] = [[] for _ in range(G)]  # nodes only role p can take in this superstep
        shared: list = []                           # nodes any role can take
        for n in remaining:
            if all(u == 0 for u in unsat[n]):
                if n in pin:
                    heapq.heappush(heaps[pin[n]], (-bl[n], n))
                else:
                    heapq.heappush(shared, (-bl[n], n))
        taken = set()
        assigned: List[int] = []
        while True:
            best = None
            for p in sorted(range(G), key=lambda q: load[q]):
                if load[p] >= thickness:
                    continue
                while heaps[p] and heaps[p][0][1] in taken:
                    heapq.heappop(heaps[p])
                while shared and shared[0][1] in taken:
                    heapq.heappop(shared)
                if heaps[p]:
                    best = (p, heapq.heappop(heaps[p])[1])
                    break
                if shared:
                    best = (p, heapq.heappop(shared)[1])
                    break
            if best is None:
                break
            p, n = best
            taken.add(n)
            proc[n], step[n] = p, s
            assigned.append(n)
            load[p] += COST[g.op[n]] + RECV_COST * sum(1 for k in preds[n] if proc[k] != p)
            for m in succ[n]:
                unsat[m][p] -= 1
                if unsat[m][p] == 0 and m not in taken and pin.get(m, p) == p:
                    heapq.heappush(heaps[p], (-bl[m], m))
        if not assigned:
            raise RuntimeError("team partition made no progress (pinned node unreachable?)")
        for n in assigned:
            p = proc[n]
            remaining.discard(n)
            for m in succ[n]:
                for q in range(G):
                    if q != p:
                        unsat[m][q] -= 1
        plan.loads.append(load)
        s += 1
    if s == 0:                                      # nothing to compute (constant outputs): one empty superstep
        plan.loads.append([0] * G)
        s = 1
    plan.nsteps = s
    _finish(plan, preds)
    return plan


def _finish(plan: TeamPlan, preds: Dict[int, List[int]]) -> None:
    """Emission order per block, exchange slots (with reuse), receive/send lists, output stores."""
    g, G, S = plan.g, plan.G, plan.nsteps
    proc, step = plan.proc, plan.step
    sched = [n for n in g.schedule(plan.outs) if n in proc]
    plan.blocks = [[[] for _ in range(G)] for _ in range(S)]
    for n in sched:
        plan.blocks[step[n]][proc[n]].append(n)
    # consumers of every published value: (role, first superstep it is needed in)
    first_need: Dict[Tuple[int, int], int] = {}
    last_need: Dict[int, int] = {}
    for n in sched:
        for k in preds[n]:
            if proc[k] != proc[n]:
                key = (k, proc[n])
                first_need[key] = min(first_need.get(key, 1 << 30), step[n])
                last_need[k] = max(last_need.get(k, -1), step[n])
    # outputs: stored by the producing role at the end of the producing superstep; constants,
    # inputs and negations of inputs are stored by role i % G in superstep 0
    plan.out_role, plan.out_step = [], []
    for i, o in enumerate(plan.outs):
        b = _base(g, o)
        if b in proc:
            plan.out_role.append(proc[b]); plan.out_step.append(step[b])
        else:
            plan.out_role.append(i % G); plan.out_step.append(0)
    # slots: a slot whose last reader works in superstep L is free for values produced in L+1
    plan.recv = [[[] for _ in range(G)] for _ in range(S)]
    plan.send = [[[] for _ in range(G)] for _ in range(S)]
    for (k, role), st in sorted(first_need.items(), key=lambda kv: (kv[1], kv[0])):
        plan.recv[st][role].append(k)
    free: List[int] = []
    release: Dict[int, List[int]] = {}
    nslots = 0
    by_step: Dict[int, List[int]] = {}
    for k in last_need:
        by_step.setdefault(step[k], []).append(k)
    for st in range(S):
        for sl in release.pop(st, []):
            free.append(sl)
        for k in sorted(by_step.get(st, [])):
            if free:
                sl = free.pop()
            else:
                sl = nslots
                nslots += 1
            plan.slot[k] = sl
            plan.send[st][proc[k]].append(k)
            release.setdefault(last_need[k] + 1, []).append(sl)
    plan.nslots = nslots


def simulate(plan: TeamPlan, inputs: Dict[Tuple[str, int], float], tables=None) -> List[float]:
    """Replay the plan: per-role private values, exchange slots, barriers.  Returns the outputs
    as the emitted code would store them (raises if a role touches a value it cannot see)."""
    g, G = plan.g, plan.G
    priv: List[Dict[int, float]] = [dict() for _ in range(G)]
    X: Dict[int, float] = {}
    owner_of_slot: Dict[int, int] = {}
    res: List[Optional[float]] = [None] * len(plan.outs)

    def val(role: int, k: int) -> float:
        o = g.op[k]
        if o == CONST:
            return g.val[k]
        if o == INPUT:
            return float(inputs[g.payload[k]])
        if o == NEG:
            return -val(role, g.a[k])
        return priv[role][k]                       # KeyError = plan bug

    for st in range(plan.nsteps):
        for role in range(G):
            for k in plan.recv[st][role]:
                assert owner_of_slot.get(plan.slot[k]) == k, "exchange slot overwritten before its last reader"
                priv[role][k] = X[plan.slot[k]]
            for n in plan.blocks[st][role]:
                sub = {}
                for k in _kids(g, n):
                    sub[k] = val(role, k)
                priv[role][n] = _eval_node(g, n, sub, tables)
        pending = []
        for role in range(G):                      # publications become visible after the barrier
            for k in plan.send[st][role]:
                pending.append((plan.slot[k], k, priv[role][k]))
        for i, o in enumerate(plan.outs):
            if plan.out_step[i] == st:
                res[i] = val(plan.out_role[i], o)
        for sl, k, v in pending:
            X[sl] = v
            owner_of_slot[sl] = k
    return res   # type: ignore


def _eval_node(g: Graph, n: int, v: Dict[int, float], tables) -> float:
    import math
    o, a, b, c = g.op[n], g.a[n], g.b[n], g.c[n]
    if o == ADD: return v[a] + v[b]
    if o == SUB: return v[a] - v[b]
    if o == MUL: return v[a] * v[b]
    if o == DIV: return v[a] / v[b]
    if o == SIN: return math.sin(v[a])
    if o == COS: return math.cos(v[a])
    if o == SQRT: return math.sqrt(v[a])
    if o == ABS: return abs(v[a])
    if o == ATAN2: return math.atan2(v[a], v[b])
    if o == MIN: return min(v[a], v[b])
    if o == MAX: return max(v[a], v[b])
    if o == SELGE: return v[b] if v[a] >= 0.0 else v[c]
    if o == INTERP:
        kind, table, deriv = g.payload[n]
        return tables.eval(kind, table, deriv, v[a])
    raise NotImplementedError(o)


# ---------------------------------------------------------------------------------------------
# emission
# ---------------------------------------------------------------------------------------------

def emit_team_function(plan: TeamPlan, name: str, signature_args: str, input_fmt: Dict[str, str],
                       out_fmt: str) -> str:
    """G role functions `name_r<w>(args)` + a dispatcher template.  `out_fmt.format(i=...)` is the
    lvalue of output i; MB_X(k) is exchange slot k; MB_TEAM_BAR() the superstep barrier.  Every
    role executes exactly `plan.nsteps` barriers."""
    g, G = plan.g, plan.G
    out: List[str] = []
    out.append(f"// team function {name}: G={G}, supersteps={plan.nsteps}, exchange slots={plan.nslots}, "
               f"work={plan.work}, makespan={plan.makespan} (model speed-up {plan.speedup:.2f})")
    for role in range(G):
        pr = Printer(g, input_fmt)
        nodes_role = [n for st in range(plan.nsteps) for n in plan.blocks[st][role]]
        pr._begin(nodes_role)
        lines = [f"MB_TEAM_FN void {name}_r{role}({signature_args}) {{"]
        for st in range(plan.nsteps):
            lines.append(f"  // ---- superstep {st}")
            for k in plan.recv[st][role]:
                lines.append(f"  const double x{k} = MB_X({plan.slot[k]});")
            for n in plan.blocks[st][role]:
                ln = pr._line(n, "  ")
                if ln is not None:
                    lines.append(ln)
            for k in plan.send[st][role]:
                lines.append(f"  MB_X({plan.slot[k]}) = x{k};")
            for i, o in enumerate(plan.outs):
                if plan.out_step[i] == st and plan.out_role[i] == role:
                    lines.append(f"  {out_fmt.format(i=i)} = {pr.ref(o)};")
            lines.append("  MB_TEAM_BAR();")
        lines.append("}")
        out.append("\n".join(lines))
    return "\n".join(out)

"""ModelBuilder: states, parameters, shared time tables and the reduced system.

A `ReducedSystem` is what `multibody(model)` + `ODEProblem` produce in the
reference (src/Multibody.jl:166-188; test/test_robot.jl:272): an ordered state
vector with initial values, tunable parameters, a diagonal mass matrix and the
right-hand side `du = f(u, p, t)` of the index-reduced system
(`prob.f(dx, x, p, t)`, src/Multibody.jl:69).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

from ..codegen.dag import Graph, Ex


@dataclass
class ReducedSystem:
    name: str
    g: Graph
    state_names: List[str]
    u0: np.ndarray                 # default initial state
    mass_diag: np.ndarray          # 1 = differential, 0 = algebraic
    param_names: List[str]
    p0: np.ndarray                 # default tunable parameters
    rhs: List[int]                 # node ids, one per state
    tables: "Tables"
    observed: Dict[str, int] = field(default_factory=dict)
    meta: Dict[str, object] = field(default_factory=dict)
    # optional hand-derived Jacobian rows {state row: {col: node}} (must equal d rhs/d u;
    # checked against plain differentiation in tests/test_codegen.py)
    jac_rows: Dict[int, Dict[int, int]] = field(default_factory=dict)

    @property
    def n(self) -> int:
        return len(self.state_names)

    @property
    def m(self) -> int:
        return len(self.param_names)

    @property
    def observed_names(self) -> List[str]:
        """Order of the generated `mbm_<model>_observed` outputs."""
        return sorted(self.observed)

    def state_index(self, name: str) -> int:
        return self.state_names.index(name)

    def param_index(self, name: str) -> int:
        return self.param_names.index(name)

    def u_nodes(self) -> List[int]:
        return [self.g.input("u", i) for i in range(self.n)]

    def t_node(self) -> int:
        return self.g.input("t", 0)

    def eval_rhs(self, u, p, t) -> np.ndarray:
        """Slow reference evaluation by DAG interpretation (tests only)."""
        inp = {("u", i): float(u[i]) for i in range(self.n)}
        inp.update({("p", j): float(p[j]) for j in range(self.m)})
        inp[("t", 0)] = float(t)
        return np.array(self.g.evaluate(self.rhs, inp, self.tables))


class Tables:
    """Uniform-grid time tables shared by all trajectories of an ensemble
    (reference: CubicSpline / LinearInterpolation with constant extrapolation,
    src/robot/path_planning.jl:191-193).

    Packed layout (f64): header [t0, dt, nk] then per table `nk` interleaved
    (value, z) pairs, z = spline second derivative (0 for linear tables)."""

    def __init__(self):
        self.t0 = 0.0
        self.dt = 1.0
        self.nk = 0
        self.values: List[np.ndarray] = []
        self.z: List[Optional[np.ndarray]] = []

    def set_grid(self, t0: float, dt: float, nk: int):
        self.t0, self.dt, self.nk = float(t0), float(dt), int(nk)

    def add(self, values: Sequence[float], cubic: bool) -> int:
        v = np.asarray(values, dtype=np.float64)
        assert v.shape == (self.nk,)
        self.values.append(v)
        self.z.append(natural_spline_z(v, self.dt) if cubic else None)
        return len(self.values) - 1

    def pack(self) -> np.ndarray:
        """[t0, dt, nk] then per table nk interleaved (value, z) pairs."""
        out = [np.array([self.t0, self.dt, float(self.nk)])]
        for v, z in zip(self.values, self.z):
            pair = np.empty((self.nk, 2))
            pair[:, 0] = v
            pair[:, 1] = z if z is not None else 0.0
            out.append(pair.ravel())
        return np.concatenate(out) if self.values else np.array([0.0, 1.0, 0.0])

    def eval(self, kind: str, table: int, deriv: int, t: float) -> float:
        """Python restatement of the device lookup (csrc/mb_tables.cuh)."""
        v = self.values[table]
        x = (t - self.t0) / self.dt
        last = self.nk - 1
        if x <= 0.0:
            return float(v[0]) if deriv == 0 else 0.0
        if x >= last:
            return float(v[last]) if deriv == 0 else 0.0
        i = min(int(x), last - 1)
        h = self.dt
        a = (x - i) * h          # t - t_i
        b = h - a                # t_{i+1} - t
        if kind == "linear":
            if deriv == 0:
                return float((v[i] * b + v[i + 1] * a) / h)
            return float((v[i + 1] - v[i]) / h) if deriv == 1 else 0.0
        z = self.z[table]
        c0 = v[i] / h - z[i] * h / 6.0
        c1 = v[i + 1] / h - z[i + 1] * h / 6.0
        if deriv == 0:
            return float(z[i] * b ** 3 / (6 * h) + z[i + 1] * a ** 3 / (6 * h) + c1 * a + c0 * b)
        if deriv == 1:
            return float(-z[i] * b ** 2 / (2 * h) + z[i + 1] * a ** 2 / (2 * h) + c1 - c0)
        if deriv == 2:
            return float((z[i] * b + z[i + 1] * a) / h)
        return float((z[i + 1] - z[i]) / h) if deriv == 3 else 0.0


def natural_spline_z(u: np.ndarray, h: float) -> np.ndarray:
    """Second derivatives of the natural cubic spline on a uniform grid
    (DataInterpolations.CubicSpline [ext]: tridiagonal system, z=0 at both ends;
    SURVEY.md Appendix B.2)."""
    n = len(u)
    z = np.zeros(n)
    if n < 3:
        return z
    # h z[i-1] + 4h z[i] + h z[i+1] = 6 (u[i+1]-2u[i]+u[i-1])/h ; Thomas algorithm
    d = 6.0 * (u[2:] - 2.0 * u[1:-1] + u[:-2]) / h
    m = n - 2
    cp = np.zeros(m)
    dp = np.zeros(m)
    cp[0] = h / (4.0 * h)
    dp[0] = d[0] / (4.0 * h)
    for i in range(1, m):
        den = 4.0 * h - h * cp[i - 1]
        cp[i] = h / den
        dp[i] = (d[i] - h * dp[i - 1]) / den
    zi = np.zeros(m)
    zi[-1] = dp[-1]
    for i in range(m - 2, -1, -1):
        zi[i] = dp[i] - cp[i] * zi[i + 1]
    z[1:-1] = zi
    return z


class ModelBuilder:
    def __init__(self, name: str):
        self.name = name
        self.g = Graph()
        self.state_names: List[str] = []
        self.u0: List[float] = []
        self.mass: List[int] = []
        self.param_names: List[str] = []
        self.p0: List[float] = []
        self.rhs: Dict[int, int] = {}
        self.tables = Tables()
        self.observed: Dict[str, int] = {}
        self.jac_rows: Dict[int, Dict[int, int]] = {}
        self._nacc = 0
        self.t = Ex(self.g, self.g.input("t", 0))

    def const(self, v: float) -> Ex:
        return Ex(self.g, self.g.const(v))

    def state(self, name: str, u0: float = 0.0, differential: bool = True) -> Ex:
        i = len(self.state_names)
        self.state_names.append(name)
        self.u0.append(float(u0))
        self.mass.append(1 if differential else 0)
        return Ex(self.g, self.g.input("u", i))

    def param(self, name: str, default: float) -> Ex:
        j = len(self.param_names)
        self.param_names.append(name)
        self.p0.append(float(default))
        return Ex(self.g, self.g.input("p", j))

    def accel(self) -> Ex:
        """Fresh unknown of a linear subsystem solved inside the RHS."""
        k = self._nacc
        self._nacc += 1
        return Ex(self.g, self.g.input("a", k))

    def index_of(self, s: Ex) -> int:
        kind, idx = self.g.payload[s.n]
        assert kind == "u"
        return idx

    def set_rhs(self, s: Ex, expr) -> None:
        e = expr if isinstance(expr, Ex) else self.const(expr)
        self.rhs[self.index_of(s)] = e.n

    def cubic(self, table: int) -> Ex:
        return Ex(self.g, self.g.interp("cubic", table, 0, self.t.n))

    def linear(self, table: int) -> Ex:
        return Ex(self.g, self.g.interp("linear", table, 0, self.t.n))

    def finish(self, **meta) -> ReducedSystem:
        n = len(self.state_names)
        missing = [self.state_names[i] for i in range(n) if i not in self.rhs]
        if missing:
            raise ValueError(f"states without an equation: {missing}")
        return ReducedSystem(self.name, self.g, list(self.state_names),
                             np.array(self.u0), np.array(self.mass, dtype=np.int32),
                             list(self.param_names), np.array(self.p0),
                             [self.rhs[i] for i in range(n)], self.tables,
                             dict(self.observed), dict(meta), dict(self.jac_rows))

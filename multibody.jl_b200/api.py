"""Host-side mirror of the reference's user API for the ensemble path.

    ssys  = multibody(model)                         # src/Multibody.jl:166-188
    prob  = ODEProblem(ssys, u0map, tspan, pmap)     # test/test_robot.jl:272-287
    eprob = EnsembleProblem(prob, prob_func=...)     # SciMLBase [ext], SURVEY.md 3.3
    sol   = solve(eprob, Rodas5P(), EnsembleB200(), trajectories=N, saveat=..., abstol=..., reltol=...)

Same names, argument meaning and error behaviour as the Julia calls; the body of
`solve(..., EnsembleB200())` is one call into libmultibody_b200 (include/mbb200.h).
Julia is absent in this image, so this Python module plays the part of
`EnsembleB200.jl` (INTEGRATION.md shows the ccall twin).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Callable, Dict, Optional, Sequence, Union

import numpy as np

from . import _lib
from .mechanics.builder import ReducedSystem
from .mechanics.zoo import MODELS, get_model


# ---- algorithms ------------------------------------------------------------------------
@dataclass(frozen=True)
class Tsit5:
    id: int = _lib.ALG_TSIT5


@dataclass(frozen=True)
class Rodas5P:
    """`autodiff` is accepted for signature parity (test/test_robot.jl:288 passes
    autodiff=false); the B200 backend always uses the generated analytic Jacobian."""
    autodiff: bool = True
    id: int = _lib.ALG_RODAS5P


@dataclass(frozen=True)
class Rodas4:
    """The solver most of the reference's tests use (test/basic_tests.jl, 25 call sites;
    docs/src/examples/ropes_and_cables.md:36 `Rodas4(autodiff=false)`)."""
    autodiff: bool = True
    id: int = _lib.ALG_RODAS4


@dataclass(frozen=True)
class FBDF:
    """The solver of the reference's loop examples (test/test_fourbar.jl:45,71 `FBDF(autodiff=true)`;
    docs/src/examples/kinematic_loops.md:80,152,190).  Served by a variable-order BDF (orders 1-5) in the published NDF
    formulation -- the same method family at the same tolerances, not OrdinaryDiffEqBDF's step sequence (DESIGN.md
    section 4).  Models rope6 / rope3e / robot6dof and the long ropes are built without this kernel
    (error MBB200_ERR_UNSUPPORTED)."""
    autodiff: bool = True
    id: int = _lib.ALG_FBDF


@dataclass
class EnsembleB200:
    """The new `SciMLBase.EnsembleAlgorithm`.  `devices`: CUDA ordinals used by this
    process (default: the first `n_gpus`)."""
    n_gpus: int = 1
    devices: Optional[Sequence[int]] = None

    def ordinals(self):
        return list(self.devices) if self.devices is not None else list(range(self.n_gpus))


# ---- model -> problem ----------------------------------------------------------------------
def multibody(model: Union[str, Callable[[], ReducedSystem], ReducedSystem]) -> ReducedSystem:
    """`multibody(model)`: returns the index-reduced system.  `model` is a zoo name, a
    builder, or an already reduced system (idempotent like `mtkcompile` on a compiled
    system)."""
    if isinstance(model, ReducedSystem):
        return model
    if isinstance(model, str):
        from .mechanics.chain import CHAIN_MODELS
        from .mechanics.zoo import IMAGE_MODELS
        if model not in MODELS and model not in CHAIN_MODELS and model not in IMAGE_MODELS:
            raise KeyError(f"unknown model {model!r}; available: {sorted(MODELS) + sorted(CHAIN_MODELS) + sorted(IMAGE_MODELS)}")
        return get_model(model)
    return model()


@dataclass
class ODEProblem:
    sys: ReducedSystem
    u0: np.ndarray
    tspan: tuple
    p: np.ndarray

    def __init__(self, sys: ReducedSystem, u0map=None, tspan=None, pmap=None):
        self.sys = sys
        self.u0 = sys.u0.copy()
        self.p = sys.p0.copy()
        self.tspan = tuple(float(x) for x in (tspan if tspan is not None else sys.meta["tspan"]))
        for k, v in (u0map or {}).items():
            if k in sys.state_names:
                self.u0[sys.state_index(k)] = float(v)
            elif k in sys.param_names:              # MTK accepts parameters in the same map
                self.p[sys.param_index(k)] = float(v)
            else:
                raise KeyError(f"{k!r} is neither a state nor a parameter of {sys.name}")
        for k, v in (pmap or {}).items():
            self.p[sys.param_index(k)] = float(v)


def remake(prob: ODEProblem, u0=None, p=None, tspan=None) -> ODEProblem:
    new = ODEProblem.__new__(ODEProblem)
    new.sys = prob.sys
    new.u0 = np.array(u0, dtype=np.float64) if u0 is not None else prob.u0.copy()
    new.p = np.array(p, dtype=np.float64) if p is not None else prob.p.copy()
    new.tspan = tuple(tspan) if tspan is not None else prob.tspan
    return new


@dataclass
class EnsembleProblem:
    """`prob_func(prob, i, repeat) -> prob` as in SciMLBase.  For large ensembles pass
    `u0_matrix` [n][N] and/or `p_matrix` [m][N] directly (the materialised result of
    running prob_func on the host)."""
    prob: ODEProblem
    prob_func: Optional[Callable] = None
    u0_matrix: Optional[np.ndarray] = None
    p_matrix: Optional[np.ndarray] = None

    def materialise(self, N: int):
        n, m = self.prob.sys.n, self.prob.sys.m
        U = np.empty((n, N))
        P = np.empty((max(m, 1), N))
        U[:] = self.prob.u0[:, None]
        P[:m] = self.prob.p[:, None] if m else 0.0
        if self.u0_matrix is not None:
            U[:] = np.asarray(self.u0_matrix, dtype=np.float64).reshape(n, N)
        if self.p_matrix is not None and m:
            P[:m] = np.asarray(self.p_matrix, dtype=np.float64).reshape(m, N)
        if self.prob_func is not None:
            for i in range(N):
                q = self.prob_func(self.prob, i, 0)
                U[:, i] = q.u0
                if m:
                    P[:m, i] = q.p
        return U, P


# ---- solution -----------------------------------------------------------------------------
@dataclass
class TrajectorySolution:
    t: np.ndarray
    u: np.ndarray            # [n_save][n]
    retcode: int
    stats: Dict[str, int]
    sys: ReducedSystem

    def __getitem__(self, name: str) -> np.ndarray:
        return self.u[:, self.sys.state_index(name)]


def successful_retcode(sol) -> bool:
    rc = sol.retcode
    return bool(np.all(np.asarray(rc) == 1))


@dataclass
class EnsembleSolution:
    sys: ReducedSystem
    t: np.ndarray            # saveat
    u: np.ndarray            # [n_save][n][N]
    t_final: np.ndarray
    retcode: np.ndarray
    stats: np.ndarray        # [5][N]: naccept, nreject, nf, njac, nsolve
    reduction: Optional[Dict[str, np.ndarray]] = None
    kernel_ms: float = 0.0
    gpu_launches: int = 0
    p: Optional[np.ndarray] = None      # [m][N] parameters of the trajectories (for observed())
    _obs: Optional[np.ndarray] = None

    def __len__(self):
        return self.u.shape[2]

    def __getitem__(self, i: int) -> TrajectorySolution:
        names = ("naccept", "nreject", "nf", "njac", "nsolve")
        return TrajectorySolution(self.t, self.u[:, :, i], int(self.retcode[i]),
                                  {k: int(self.stats[j, i]) for j, k in enumerate(names)}, self.sys)

    def observed(self, names=None) -> Dict[str, np.ndarray]:
        """`sol(t, idxs = expr)` for eliminated variables (test/test_robot.jl:321-325): evaluates
        the model's observed expressions on the device for every saved state.  Returns
        {name: [n_save][N]} (all observed variables, or the requested ones)."""
        sys = self.sys
        all_names = sys.observed_names
        if not all_names:
            raise KeyError(f"{sys.name} defines no observed variables")
        want = list(all_names) if names is None else ([names] if isinstance(names, str) else list(names))
        for nm in want:
            if nm not in all_names:
                raise KeyError(f"{nm!r} is not an observed variable of {sys.name}; available: {all_names}")
        if self._obs is None:
            ns, n, N = self.u.shape
            out = np.empty((ns, len(all_names), N))
            h = _handles[sys.name]
            P = np.ascontiguousarray(self.p) if sys.m else None
            _lib.check(_lib.lib().mbb200_observed(h.handle, N, _lib.ptr(np.ascontiguousarray(self.u)),
                                                  _lib.ptr(P) if sys.m else None, _lib.ptr(np.ascontiguousarray(self.t)),
                                                  ns, _lib.ptr(out)))
            self._obs = out
        return {nm: self._obs[:, all_names.index(nm), :] for nm in want}

    def state(self, name: str) -> np.ndarray:
        """[n_save][N] values of one named state."""
        return self.u[:, self.sys.state_index(name), :]


_handles: Dict[str, _lib.ModelHandle] = {}
_devices = None


def _ensure(ens: EnsembleB200, sys: ReducedSystem) -> _lib.ModelHandle:
    global _devices
    ords = ens.ordinals()
    if _devices is None:
        _lib.init_devices(ords)
        _devices = ords
    elif _devices != ords:
        raise _lib.Mbb200Error(-2, f"library already initialised on devices {_devices}; got {ords}")
    if sys.name not in _handles:
        try:        # fast path: a kernel set compiled into the library ahead of time
            _handles[sys.name] = _lib.ModelHandle(sys.name, sys.n, sys.m, sys.tables.pack())
        except _lib.Mbb200Error as e:
            if e.code != -3 or sys.meta.get("chain"):       # MBB200_ERR_UNKNOWN_MODEL
                raise
            # any other reduced system: the image carries the model and the library compiles it at load time (NVRTC)
            from .codegen.mbir import serialize
            _handles[sys.name] = _lib.ModelHandle(sys.name, sys.n, sys.m, sys.tables.pack(), image=serialize(sys))
    return _handles[sys.name]


def shutdown():
    global _devices
    for h in _handles.values():
        h.free()
    _handles.clear()
    if _devices is not None:
        _lib.lib().mbb200_shutdown()
        _devices = None


def make_opts(alg, tspan, abstol=1e-6, reltol=1e-3, dt=0.0, dtmin=0.0, dtmax=0.0, maxiters=0, reduce=False,
              qsteady_min=0.0, qsteady_max=0.0, exact_pow=False):
    """`qsteady_min/max` are OrdinaryDiffEq's solve keywords of the same name (0, 0 = the method's
    default: (1, 1) Tsit5, (1, 6/5) Rosenbrock); `exact_pow` replaces the controller's `fastpower`."""
    return _lib.Opts(alg.id, _lib.FLAG_REDUCE if reduce else 0, tspan[0], tspan[1], abstol, reltol,
                     dt, dtmin, dtmax, int(maxiters), qsteady_min, qsteady_max, 1 if exact_pow else 0, 0)


def solve(eprob: Union[EnsembleProblem, ODEProblem], alg, ensemblealg: Optional[EnsembleB200] = None, *,
          trajectories: int = 1, saveat=None, abstol: float = 1e-6, reltol: float = 1e-3, dt: float = 0.0,
          dtmin: float = 0.0, dtmax: float = 0.0, maxiters: int = 0, reduction: bool = False,
          qsteady_min: float = 0.0, qsteady_max: float = 0.0, exact_pow: bool = False) -> EnsembleSolution:
    """`solve(eprob, alg, EnsembleB200(); trajectories, saveat, abstol, reltol, ...)`.
    Defaults are SciML's (abstol 1e-6, reltol 1e-3).  A plain ODEProblem is run as an
    ensemble of one.  `reduction=True` also returns count/mean/variance statistics."""
    if isinstance(eprob, ODEProblem):
        eprob = EnsembleProblem(eprob)
    ens = ensemblealg or EnsembleB200()
    sys = eprob.prob.sys
    N = int(trajectories)
    h = _ensure(ens, sys)
    U, P = eprob.materialise(N)
    tspan = eprob.prob.tspan
    sv = np.ascontiguousarray(np.atleast_1d(np.asarray(saveat if saveat is not None else [tspan[1]], dtype=np.float64)))
    ns, n = len(sv), sys.n
    out = np.empty((ns, n, N))
    tf = np.empty(N)
    rc = np.zeros(N, np.int32)
    st = np.zeros((5, N), np.int32)
    red = np.zeros(2 * ns * n + 1) if reduction else None
    o = make_opts(alg, tspan, abstol, reltol, dt, dtmin, dtmax, maxiters, reduction, qsteady_min, qsteady_max, exact_pow)
    _lib.check(_lib.lib().mbb200_solve(h.handle, C.byref(o), N, _lib.ptr(U), _lib.ptr(P) if sys.m else None,
                                       _lib.ptr(sv), ns, _lib.ptr(out), _lib.ptr(tf), _lib.ptr(rc), _lib.ptr(st),
                                       _lib.ptr(red) if reduction else None))
    ms, nl = 0.0, 0
    for g in range(len(ens.ordinals())):
        m_, l_ = _lib.last_timing(g)
        ms, nl = max(ms, m_), nl + l_
    redd = None
    if reduction:
        from .distributed import finalize_statistics
        redd = finalize_statistics(red, ns, n)   # already summed over the ranks of the library communicator, if any
    return EnsembleSolution(sys, sv, out, tf, rc, st, redd, ms, nl, P[:sys.m] if sys.m else None)

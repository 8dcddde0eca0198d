"""The BASELINE.json configs as concrete synthetic ensembles.

All draws come from a counter-based generator (splitmix64 keyed by seed, trajectory index
and parameter slot), so any shard of any size reproduces the same values on every rank
and on the CPU oracle without storing or communicating them.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

from .mechanics.zoo import get_model

SEED = 20261104


def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x.copy()
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def uniform(index: np.ndarray, slot: int, lo: float, hi: float, seed: int = SEED) -> np.ndarray:
    """U[lo,hi) for trajectory indices `index` (int64 array) and parameter slot `slot`."""
    with np.errstate(over="ignore"):
        key = _splitmix64(np.uint64(seed) + np.uint64(0x632BE59BD9B4E019) * np.uint64(slot + 1))
        z = _splitmix64(index.astype(np.uint64) * np.uint64(0xD1342543DE82EF95) + key)
    return lo + (hi - lo) * ((z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0))


@dataclass
class EnsembleSpec:
    config: str
    model: str
    alg: str                 # "tsit5" | "rodas5p"
    tspan: tuple
    saveat: np.ndarray
    abstol: float
    reltol: float
    u0: np.ndarray           # [n][N]
    p: np.ndarray            # [m][N]
    n_total: int             # size the BASELINE config names


def make_ensemble(config: str, N: Optional[int] = None, offset: int = 0, links: Optional[int] = None) -> EnsembleSpec:
    """Trajectories [offset, offset+N) of config C1..C5 (N defaults to the full size).
    `links` (20 or 50) selects the BASELINE-size ropes of config 5: flexible for "C5"/"C5E" (c = 800, d = 0.01),
    stiff for "C5S"; without it C5 / C5E are the documented 6-link / 3-link ropes on the generated code path."""
    config = config.upper()
    if links is not None:
        if config not in ("C5", "C5E", "C5S") or links not in (20, 50):
            raise KeyError("links = 20 or 50 applies to the rope configs C5 / C5E (flexible) and C5S (stiff)")
        config = f"C5S{links}" if config == "C5S" else f"C5L{links}"
    if config == "C1":      # pendulum, 1024 initial angles, Tsit5 (docs/src/examples/pendulum.md:25-68)
        total, model, alg, tol = 1024, "pendulum", "tsit5", 1e-8
        saveat = np.linspace(0.0, 10.0, 101)
    elif config == "C2":    # spring-damper, 1M parameter sets, Tsit5, 100 saves (spring_damper_system.md:21-63)
        total, model, alg, tol = 1 << 20, "spring_damper", "tsit5", 1e-8
        saveat = 0.05 * np.arange(1, 101)
    elif config == "C2B":   # three-springs (massless node -> DAE), 1M parameter sets, Rodas5P (three_springs.md:22-44)
        total, model, alg, tol = 1 << 20, "three_springs", "rodas5p", 1e-8
        saveat = 0.1 * np.arange(1, 101)
    elif config == "C3":    # fourbar closed loop, 256k crank speeds, Rodas5P (test/test_fourbar.jl:13-46)
        total, model, alg, tol = 1 << 18, "fourbar", "rodas5p", 1e-8
        saveat = np.linspace(0.0, 1.4399, 21)[1:]
    elif config == "C3P":   # planar loop with RevolutePlanarLoopConstraint (kinematic_loops.md:38-80), initial crank angles
        total, model, alg, tol = 1 << 16, "fourbar_planar", "rodas5p", 1e-8
        saveat = np.linspace(0.0, 0.6, 21)[1:]
    elif config == "C4":    # Robot6DOF parameter sweep, 1M, Rodas5P reltol 1e-8 (test/test_robot.jl:268-288)
        total, model, alg, tol = 1 << 20, "robot6dof", "rodas5p", 1e-8
        saveat = np.linspace(0.0, 2.0, 21)
    elif config == "C5":    # stiff 6-link rope of docs/src/examples/ropes_and_cables.md:24-36, 64k parameter sets
        total, model, alg, tol = 1 << 16, "rope6", "rodas5p", 1e-6
        saveat = np.linspace(0.0, 5.0, 51)[1:]
    elif config == "C5E":   # flexible rope (c = 800, d = 0.01; ropes_and_cables.md:48-60) at the 3-link size of test/basic_tests.jl:928
        total, model, alg, tol = 1 << 16, "rope3e", "rodas5p", 1e-6
        saveat = np.linspace(0.0, 5.0, 51)[1:]
    elif config.startswith("C5L") or config.startswith("C5S"):
        # BASELINE.json config 5: "rope with N = 20-50 links ... warp-per-trajectory Rodas5P, 64k trajectories"; SURVEY 8d:
        # elastic c = 800, d = 0.01, tspan (0, 5), tolerance 1e-6, 50 save points
        nl = int(config[3:])
        total, model, alg, tol = 1 << 16, (f"rope{nl}e" if config[2] == "L" else f"rope{nl}"), "rodas5p", 1e-6
        saveat = np.linspace(0.0, 5.0, 51)[1:]
    else:
        raise KeyError(f"config {config} is not built (C1..C5 are)")
    N = total if N is None else int(N)
    sys = get_model(model)
    idx = np.arange(offset, offset + N, dtype=np.int64)
    U = np.repeat(sys.u0[:, None], N, axis=1)
    P = np.repeat(sys.p0[:, None], N, axis=1) if sys.m else np.zeros((0, N))
    if config == "C1":
        U[sys.state_index("joint.phi")] = -np.pi + 2.0 * np.pi * (idx + 0.5) / 1024.0
    elif config == "C2":
        c = uniform(idx, 0, 10.0, 100.0)
        d = uniform(idx, 1, 0.5, 5.0)
        m = uniform(idx, 2, 0.5, 2.0)
        P[sys.param_index("spring1.c")] = c
        P[sys.param_index("spring2.c")] = c
        P[sys.param_index("damper1.d")] = d
        P[sys.param_index("body1.m")] = m
        P[sys.param_index("body2.m")] = m
    elif config == "C2B":
        P[sys.param_index("spring1.c")] = uniform(idx, 0, 10.0, 40.0)
        P[sys.param_index("spring3.c")] = uniform(idx, 1, 10.0, 40.0)
        P[sys.param_index("spring2.c")] = uniform(idx, 2, 20.0, 80.0)
        P[sys.param_index("body1.m")] = uniform(idx, 3, 0.5, 2.0)
    elif config == "C3":
        U[sys.state_index("j1.w")] = uniform(idx, 0, 3.0, 7.0)
    elif config == "C3P":
        U[sys.state_index("j1.phi")] = uniform(idx, 0, 0.05, 0.2)     # dependents are re-initialised on the device
        P[sys.param_index("damper1.d")] = uniform(idx, 1, 0.05, 0.5)
        P[sys.param_index("damper2.d")] = uniform(idx, 2, 0.05, 0.5)
    elif config == "C5":
        P[sys.param_index("rope.d_joint")] = uniform(idx, 0, 0.5, 2.0)
        P[sys.param_index("body.m")] = uniform(idx, 1, 0.5, 2.0)
        P[sys.param_index("rope.air_resistance")] = uniform(idx, 2, 0.0, 0.4)
    elif config == "C5E":
        P[sys.param_index("rope.d")] = 0.01
        P[sys.param_index("rope.d_joint")] = uniform(idx, 0, 0.05, 0.2)
        P[sys.param_index("body.m")] = uniform(idx, 1, 100.0, 500.0)
        P[sys.param_index("rope.air_resistance")] = uniform(idx, 2, 0.0, 0.4)
    elif config.startswith("C5L"):      # d_joint U[0.05,0.2], body mass U[100,500], air_resistance U[0,0.4]
        P[sys.param_index("rope.c")] = 800.0
        P[sys.param_index("rope.d")] = 0.01
        P[sys.param_index("rope.d_joint")] = uniform(idx, 0, 0.05, 0.2)
        P[sys.param_index("body.m")] = uniform(idx, 1, 100.0, 500.0)
        P[sys.param_index("rope.air_resistance")] = uniform(idx, 2, 0.0, 0.4)
    elif config.startswith("C5S"):      # the stiff docs rope (ropes_and_cables.md:24-36) at BASELINE length, ranges of C5
        P[sys.param_index("rope.d_joint")] = uniform(idx, 0, 0.5, 2.0)
        P[sys.param_index("body.m")] = uniform(idx, 1, 0.5, 2.0)
        P[sys.param_index("rope.air_resistance")] = uniform(idx, 2, 0.0, 0.4)
    elif config == "C4":
        P[sys.param_index("mechanics.load.m")] = uniform(idx, 0, 5.0, 25.0)
        for a in range(6):
            P[sys.param_index(f"axis{a+1}.controller.kp")] = uniform(idx, 1 + 3 * a, 3.0, 8.0)
            P[sys.param_index(f"axis{a+1}.controller.ks")] = uniform(idx, 2 + 3 * a, 0.3, 0.8)
            P[sys.param_index(f"axis{a+1}.controller.Ts")] = uniform(idx, 3 + 3 * a, 0.03, 0.08)
    # C3P stops short of the flat (singular) configuration of the parallelogram
    tspan = (0.0, 0.6) if config == "C3P" else tuple(sys.meta["tspan"])
    return EnsembleSpec(config, model, alg, tspan, saveat, tol, tol,
                        np.ascontiguousarray(U), np.ascontiguousarray(P), total)

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


MODELS: Dict[str, Callable[[], ReducedSystem]] = {
    "pendulum": pendulum,
    "spring_damper": spring_damper,
    "robot6dof": robot6dof,
}
_cache: Dict[str, ReducedSystem] = {}


def get_model(name: str) -> ReducedSystem:
    if name not in _cache:
        _cache[name] = MODELS[name]()
    return _cache[name]

"""Long `Rope` models (BASELINE.json config 5: 20-50 links).

`Rope` is reference src/components.jl:574-626: n+1 `Spherical(state=true)` joints and n links, stiff
(`BodyShape`) or flexible (`Prismatic` + `Translational.Spring`/`Damper` + `Body`, :594-611).  The
straight-line code generator is O(n^3) in code size on a chain, so these sizes do not go through the
expression DAG: their right-hand side is the hand-written O(N) articulated-body sweep of
`csrc/mb_chain.h` (one source for the CUDA kernels and the CPU oracle), and their Jacobian is differenced
forward on the device, as the reference's rope examples do (`Rodas4(autodiff=false)`,
docs/src/examples/ropes_and_cables.md:36,61).  State and parameter names, order and defaults are those of
the generated models `rope6` / `rope3e` (mechanics/zoo.py); `rope6c` / `rope3ec` are the loop-based code at
those two sizes, which is how it is pinned (tests/test_chain.py).  The list below mirrors
`csrc/mb_chain_models.h`.
"""
from __future__ import annotations

from typing import Callable, Dict

import numpy as np

from ..codegen.dag import Graph
from .builder import ReducedSystem, Tables

CHAIN_SPECS = {"rope6c": (6, False), "rope3ec": (3, True), "rope20": (20, False), "rope50": (50, False),
               "rope20e": (20, True), "rope50e": (50, True)}


def chain_rope(name: str) -> ReducedSystem:
    n_links, elastic = CHAIN_SPECS[name]
    li = 1.0 / n_links
    names, u0 = [], []
    for i in range(n_links + 1):
        names += [f"rope.joint_{i+1}.phi[{k+1}]" for k in range(3)] + [f"rope.joint_{i+1}.phid[{k+1}]" for k in range(3)]
        u0 += [0.0] * 6
        if elastic and i < n_links:
            names += [f"rope.flexibility_{i+1}.s", f"rope.flexibility_{i+1}.v"]
            u0 += [li, 0.0]
    if elastic:     # defaults of zoo.rope_elastic (test/basic_tests.jl:928-959)
        pn, p0 = ["rope.d_joint", "body.m", "rope.air_resistance", "rope.c", "rope.d"], [0.1, 300.0, 0.2, 800.0, 0.001]
    else:           # defaults of zoo.rope (ropes_and_cables.md:24-36)
        pn, p0 = ["rope.d_joint", "body.m", "rope.air_resistance"], [1.0, 1.0, 0.0]
    return ReducedSystem(name=name, g=Graph(), state_names=names, u0=np.array(u0), mass_diag=np.ones(len(names), dtype=int),
                         param_names=pn, p0=np.array(p0), rhs=[], tables=Tables(),
                         meta={"tspan": (0.0, 5.0), "chain": (n_links, elastic), "jacobian": "forward differences on the device"})


CHAIN_MODELS: Dict[str, Callable[[], ReducedSystem]] = {k: (lambda k=k: chain_rope(k)) for k in CHAIN_SPECS}

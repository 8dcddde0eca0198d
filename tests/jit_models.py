"""Models that are NOT compiled into libmultibody_b200: they reach the device only through an MBIR2 image that carries
them (codegen/mbir.py -> mbb200_model_load -> NVRTC)."""
import numpy as np

from multibody_jl_b200.mechanics.builder import ModelBuilder
from multibody_jl_b200.mechanics.mechanism import Mechanism


def double_pendulum():
    """test/basic_tests.jl:274-325: two revolute joints (n = normalize([0.1, 0, 1]) and [0, 0, 1]) with
    `Rotational.Damper`s d = 5 and d = 1 on their axes, rods `FixedTranslation(r = [1, 0, 0])`, point masses m = 1;
    started at phi = (1, 0.5); the reference asserts rev1.phi(35) ~ -pi/2 and body2.r_0[2](35) ~ -2 (rtol 1e-2)."""
    mb = ModelBuilder("double_pendulum_jit")
    mech = Mechanism(mb)
    d1 = mb.param("damper1.d", 5.0)
    d2 = mb.param("damper2.d", 1.0)
    p1 = mb.state("rev1.phi", 1.0)
    w1 = mb.state("rev1.w", 0.0)
    p2 = mb.state("rev2.phi", 0.5)
    w2 = mb.state("rev2.w", 0.0)
    n1 = np.array([0.1, 0.0, 1.0]) / np.linalg.norm([0.1, 0.0, 1.0])
    j1 = mech.revolute(mech.world(), tuple(n1), p1, w1, tau_applied=-d1 * w1, name="rev1")
    a = mech.fixed_translation(j1, (1.0, 0.0, 0.0))
    mech.body(a, 1.0, (0.0, 0.0, 0.0))
    j2 = mech.revolute(a, (0.0, 0.0, 1.0), p2, w2, tau_applied=-d2 * w2, name="rev2")
    b = mech.fixed_translation(j2, (1.0, 0.0, 0.0))
    mech.body(b, 1.0, (0.0, 0.0, 0.0))
    for c in range(3):
        v = b.r0[c]
        mb.observed[f"body2.r_0[{c+1}]"] = v.n if hasattr(v, "n") else mb.const(float(v)).n
    mech.solve_accelerations()
    return mb.finish(tspan=(0.0, 35.0), mechanism=mech)

"""`FBDF()` (the solver of the reference's loop examples, test/test_fourbar.jl:45,71) as a variable-order BDF in the NDF
formulation (oracle/mbo_bdf.inc; the device kernel csrc/mb_bdf.cuh is the same arithmetic).  The reference's own FBDF is
un-vendored, so nothing here is a step-by-step pin: the oracle is checked against the independent golden vectors, against
the Rosenbrock methods, and for the convergence a multistep method of variable order must show."""
import os

import numpy as np
import pytest

from multibody_jl_b200 import ensembles
from multibody_jl_b200.mechanics.zoo import get_model

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _solve(oracle, spec, alg, n=None, **kw):
    sysm = get_model(spec.model)
    kw.setdefault("abstol", spec.abstol)
    kw.setdefault("reltol", spec.reltol)
    u0, p = (spec.u0, spec.p) if n is None else (spec.u0[:, :n], spec.p[:, :n])
    return oracle.solve(spec.model, alg, u0, p, sysm.tables.pack(), spec.tspan, spec.saveat, **kw)


def _err(a, b):
    scale = np.maximum(np.abs(b).max(axis=(0, 2), keepdims=True), 1e-3)
    return (np.abs(a - b) / scale).max()


@pytest.mark.parametrize("config,tol_loose,tol_tight", [("C1", 5e-4, 2e-5), ("C3", 2e-3, 5e-5), ("C3P", 1e-4, 5e-6),
                                                        ("C2B", 2e-3, 1e-4), ("C4", 5e-5, 5e-6)])
def test_bdf_converges_to_the_rosenbrock_solution(oracle, config, tol_loose, tol_tight):
    """ODEs, index-1 DAEs (both loop models of test/test_fourbar.jl) and the robot: every trajectory succeeds and the
    global error falls with the tolerance (tol 1e-6 -> 1e-8: at least x8)."""
    n = 4 if config == "C4" else 16
    spec = ensembles.make_ensemble(config, n)
    ref = _solve(oracle, spec, "rodas5p", abstol=1e-11, reltol=1e-11)
    assert (ref["retcode"] == 1).all()
    r6 = _solve(oracle, spec, "fbdf", abstol=1e-6, reltol=1e-6)
    r8 = _solve(oracle, spec, "fbdf", abstol=1e-8, reltol=1e-8)
    assert (r6["retcode"] == 1).all() and (r8["retcode"] == 1).all()
    e6, e8 = _err(r6["u"], ref["u"]), _err(r8["u"], ref["u"])
    print(config, "BDF global error: tol 1e-6 ->", e6, " 1e-8 ->", e8, " accepted steps", r6["stats"][0].mean(), r8["stats"][0].mean())
    assert e6 < tol_loose and e8 < tol_tight and e8 < e6 / 8
    assert r8["stats"][0].mean() > r6["stats"][0].mean()


def test_bdf_fourbar_against_the_independent_lagrange_golden(oracle):
    """The reference's FBDF call site itself (test/test_fourbar.jl:44-46: solve(prob, FBDF(autodiff=true)) on (0, 1.4399),
    `@test SciMLBase.successful_retcode(sol)`), checked by value against the 1-DOF Lagrange derivation."""
    gold = np.load(os.path.join(GOLD, "fourbar_lagrange.npz"))
    sysm = get_model("fourbar")
    for k, w0 in enumerate(gold["w0"]):
        u0 = sysm.u0.copy()
        u0[sysm.state_index("j1.w")] = w0
        r = oracle.solve("fourbar", "fbdf", u0[:, None], np.zeros((1, 1)), sysm.tables.pack(), (0.0, 1.4399),
                         gold["saveat"], abstol=1e-9, reltol=1e-9)
        assert r["retcode"][0] == 1
        assert np.abs(r["u"][:, :, 0] - gold["u"][k]).max() < 2e-4
        for j in (0, 7, 19):                                      # the loop stays closed at the interpolated points
            assert np.abs(sysm.eval_rhs(r["u"][j, :, 0], sysm.p0, 0.0)[2:]).max() < 1e-6


def test_bdf_uses_high_orders_and_few_solves_on_smooth_problems(oracle):
    """A multistep method pays ~2 right-hand sides per step where Rodas5P pays 8 plus a Jacobian: on the smooth planar
    loop the BDF run needs fewer linear solves per unit time than Rodas5P needs stage solves at the same tolerance
    only if the order climbs -- which is what this asserts (steps at tol 1e-8 would be ~20x more at order 1-2)."""
    spec = ensembles.make_ensemble("C3P", 8)
    rb = _solve(oracle, spec, "fbdf")
    rr = _solve(oracle, spec, "rodas5p")
    assert (rb["retcode"] == 1).all()
    nacc, nsolve, nf = rb["stats"][0].mean(), rb["stats"][4].mean(), rb["stats"][2].mean()
    print("BDF naccept", nacc, "nsolve", nsolve, "nf", nf, "| Rodas5P naccept", rr["stats"][0].mean(), "nsolve", rr["stats"][4].mean())
    assert nsolve / nacc < 3.0                                    # Newton converges in 1-3 iterations
    assert nacc < 60 * rr["stats"][0].mean()                      # order >= 4 most of the way (order 2 would be > 200x)


def test_bdf_dense_output_and_failure_codes(oracle):
    spec = ensembles.make_ensemble("C1", 4)
    a = _solve(oracle, spec, "fbdf")
    sv = spec.saveat
    spec.saveat = sv[::3]                                         # dense output does not change the steps taken
    b = _solve(oracle, spec, "fbdf")
    assert np.array_equal(a["u"][::3], b["u"]) and np.array_equal(a["stats"], b["stats"])
    r = _solve(oracle, spec, "fbdf", maxiters=5)
    assert (r["retcode"] == 4).all() and np.isnan(r["u"][-1]).all()
    spec = ensembles.make_ensemble("C3", 4)                       # inconsistent algebraic states that Newton cannot repair
    spec.u0[2:] = 1e3
    r = _solve(oracle, spec, "fbdf")
    assert (r["retcode"] == 7).all()                             # ReturnCode.InitialFailure, as the Rosenbrock path

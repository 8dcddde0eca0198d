"""Warp-team partition (codegen/team.py): every plan is replayed numerically -- role-private
values, exchange slots with reuse, barriers -- and must reproduce the plain DAG evaluation."""
import numpy as np
import pytest

from multibody_jl_b200.codegen import team
from multibody_jl_b200.codegen.model_emit import ModelCode
from multibody_jl_b200.mechanics.zoo import get_model


def _inputs(sysm, rng, extra=()):
    u = np.array(sysm.u0, float) + 0.1 * rng.standard_normal(sysm.n)
    p = np.array(sysm.p0, float) * (1.0 + 0.05 * rng.standard_normal(sysm.m))
    inp = {("u", i): u[i] for i in range(sysm.n)}
    inp.update({("p", i): p[i] for i in range(sysm.m)})
    inp[("t", 0)] = 0.7345
    inp[("c", 0)] = 1.0 / (0.21193756319429014 * 1.3e-3)
    for k, v in extra:
        inp[k] = v
    return inp


@pytest.mark.parametrize("model,G", [("robot6dof", 4), ("robot6dof", 8), ("furuta", 2)])
def test_team_plans_reproduce_the_dag(model, G):
    sysm = get_model(model)
    code = ModelCode(sysm, pivots=sysm.meta.get("pivots"))
    g = sysm.g
    rng = np.random.default_rng(5)
    inp = _inputs(sysm, rng)
    for name, outs in (("rhs", list(sysm.rhs)), ("factor", list(code.lu_nodes)), ("dT", list(code.dT_nodes))):
        plan = team.partition(g, outs, G)
        got = team.simulate(plan, inp, sysm.tables)
        ref = g.evaluate(outs, inp, sysm.tables)
        assert np.allclose(got, ref, rtol=0, atol=0), name          # same operations, same order: bit equal
        assert sum(len(b) for st in plan.blocks for b in st) == len(plan.proc)
        assert plan.nslots <= len(plan.slot)
    # solve lives on its own small graph
    sg = code.solve_graph
    sinp = {("lu", i): rng.standard_normal() for i in range(max(code.nlu, 1))}
    sinp.update({("b", i): rng.standard_normal() for i in range(sysm.n)})
    sinp.update({("p", i): inp[("p", i)] for i in range(sysm.m)})
    sinp[("c", 0)] = inp[("c", 0)]
    plan = team.partition(sg, list(code.solve_out), G)
    assert team.simulate(plan, sinp, None) == sg.evaluate(list(code.solve_out), sinp, None)


def test_team_partition_balances_the_robot_rhs():
    sysm = get_model("robot6dof")
    plan = team.partition(sysm.g, list(sysm.rhs), 4)
    assert plan.speedup > 1.4 and plan.nsteps <= 20        # model figures; the GPU numbers are in profiles/

"""Step-control variants of the oracle against the reference-run 16-digit outputs.

Reference vectors: test/test_PlanarMechanics.jl:59 (planar pendulum, Rodas5P default tol),
test/test_orientation_getters.jl:69 (Furuta, Rodas5P default tol), test/test_worldforces.jl:136,180
(Tsit5, reltol 1e-8).  Prints |oracle - printed value| per variant (committed: profiles/r02_pin_table.md)."""
import itertools, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import _loadpkg                     # noqa
_loadpkg.load()
from multibody_jl_b200.mechanics.zoo import get_model
from oracle import mbo
KA = json.load(open(os.path.join(ROOT, "tests/golden/reference_known_answers.json")))


def planar(**kw):
    s = get_model("planar_pendulum"); case = KA["planar_pendulum_rodas5p_default_tolerances"]
    r = mbo.solve("planar_pendulum", "rodas5p", np.array(s.u0, float)[:, None], np.zeros((0, 1)), s.tables.pack(), (0.0, 3.0),
                  np.array([1.0]), abstol=case["abstol"], reltol=case["reltol"], **kw)
    return abs(r["u"][0, 0, 0] - case["rod_frame_a_phi_t1"]), r["stats"][:2, 0]


def furuta(**kw):
    s = get_model("furuta"); case = KA["furuta_rodas5p_default_tolerances"]
    tb = s.tables.pack(); p = np.array(s.p0, float)[:, None]
    r = mbo.solve("furuta", "rodas5p", np.array(s.u0, float)[:, None], p, tb, (0.0, 12.0), np.array([1.0, 12.0]),
                  abstol=case["abstol"], reltol=case["reltol"], **kw)
    names = s.observed_names
    ua = [names.index(f"upper_arm.frame_b.r_0[{c}]") for c in (1, 2, 3)]
    la = [names.index(f"lower_arm.frame_b.r_0[{c}]") for c in (1, 2, 3)]
    o1 = mbo.eval_fn("furuta", 3, np.ascontiguousarray(r["u"][0, :, 0]), p[:, 0], 1.0, tb)
    o12 = mbo.eval_fn("furuta", 3, np.ascontiguousarray(r["u"][1, :, 0]), p[:, 0], 12.0, tb)
    return (np.abs(o1[ua] - case["upper_arm_frame_b_t1"]).max(), np.abs(o12[la] - case["lower_arm_frame_b_t12"]).max()), r["stats"][:2, 0]


def couple(key, **kw):
    s = get_model("free_couple"); case = KA[key]
    p = np.array(s.p0, float); p[s.param_index("resolve_world")] = case["resolve_world"]
    r = mbo.solve("free_couple", "tsit5", np.array(s.u0, float)[:, None], p[:, None], s.tables.pack(), (0.0, case["t"]),
                  np.array([case["t"]]), abstol=case["abstol"], reltol=case["reltol"], **kw)
    return np.abs(r["u"][0, :3, 0] - np.array(case["frame_a_r_0"])).max(), r["stats"][:2, 0]


if __name__ == "__main__":
    print("| dead band | pow | planar phi(1) | furuta r(1) | furuta r(12) | couple world | couple body |")
    print("|---|---|---|---|---|---|---|")
    for qs, pm in itertools.product(((1.0, 1.0), (1.0, 1.2)), (1, 2)):
        a, sa = planar(qsteady=qs, pow_mode=pm)
        (b1, b12), sb = furuta(qsteady=qs, pow_mode=pm)
        c, sc = couple("worldforce_couple_world_frame", qsteady=(1.0, 1.0), pow_mode=pm)
        d, sd = couple("worldforce_couple_body_frame", qsteady=(1.0, 1.0), pow_mode=pm)
        print(f"| {qs} | {'exact' if pm == 1 else 'fastpower'} | {a:.2e} ({sa[0]}+{sa[1]}) | {b1:.2e} ({sb[0]}+{sb[1]}) | {b12:.2e} | {c:.2e} ({sc[0]}+{sc[1]}) | {d:.2e} ({sd[0]}+{sd[1]}) |")

"""Phase cycle counts of the long-rope warp kernel (library built with -DMBC_PROFILE, selected with MBB200_LIB)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mbpkg; mbpkg.load()
import numpy as np
from multibody_jl_b200 import api, ensembles
for cfg, links, N, t1 in (("C5S", 20, 64, 0.5), ("C5", 20, 64, 0.1), ("C5S", 50, 64, 0.25)):
    spec = ensembles.make_ensemble(cfg, N, links=links)
    s = api.multibody(spec.model)
    ep = api.EnsembleProblem(api.ODEProblem(s, tspan=(0.0, t1)), u0_matrix=spec.u0, p_matrix=spec.p)
    t = time.time()
    sol = api.solve(ep, api.Rodas5P(), api.EnsembleB200(), trajectories=N, saveat=np.array([t1]), abstol=1e-6, reltol=1e-6)
    print(cfg, links, "kernel ms %.0f" % sol.kernel_ms, "steps", sol.stats[0].mean() + sol.stats[1].mean(), flush=True)

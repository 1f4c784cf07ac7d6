"""DEVELOPMENT TOOL: accuracy of knock-down end states on the GPU vs LSODA 1e-13."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle"))
import lassim_oracle as orc
from lassim_b200._capi import Engine
gold = os.path.join(REPO, "tests", "golden")
p = dict(np.load(os.path.join(gold, "th2_problem.npz")))
g = dict(np.load(os.path.join(gold, "th2_uniform.npz")))
n = 12
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
X = g["X"][:6]
r = eng.eval(X, details=True)
E = eng.last_end_states(len(X))
ode = orc.tight_ode()
for c, x in enumerate(X):
    sol = x.copy()
    ctrl = {}
    worst = 0; worst_rel = 0
    for i in range(n):
        tt = p["pert"][i, n:]
        key = str(tt)
        if key not in ctrl:
            ctrl[key] = ode(p["y0"], tt, sol, p["vector_map"], p["mask"], n)[-1]
            u = len(ctrl)
            d = np.abs(E[c, u] - ctrl[key]); s = np.maximum(np.abs(ctrl[key]), 1e-12)
            print("cand", c, "ctrl", u, "max abs err %.2e max rel(floor 1e-12) %.2e  min|ctrl| %.2e" % (d.max(), (d / s).max(), np.abs(ctrl[key]).min()))
        sol[i] *= p["pert"][i, i]
        e = ode(p["y0"], tt, sol, p["vector_map"], p["mask"], n)[-1]
        d = np.abs(E[c, 3 + i] - e)
        rel = d / np.maximum(np.abs(e), 1e-8)
        worst = max(worst, d.max()); worst_rel = max(worst_rel, rel.max())
    print("cand", c, "knock worst abs %.2e worst rel(floor 1e-8) %.2e ; pert gpu %.6f tight %.6f ref %.6f" % (worst, worst_rel, r["pert_cost"][c], g["tight_pert_cost"][c], g["ref_pert_cost"][c]))
    # reduction check
    sim = np.empty(n * n)
    for i in range(n):
        u = 1 if p["pert"][i, n + 1] == 24 else 2
        sim[i * n:(i + 1) * n] = E[c, 3 + i] / E[c, u]
    sim -= 1; sim[sim > 2] = 2
    print("   lstsq on gpu E:", orc.rank1_residual(sim, p["pert"][:, :n].flatten()), "gpu", r["pert_cost"][c])

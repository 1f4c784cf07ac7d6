"""DEVELOPMENT TOOL: knock-down end states of selected rows of the seed-0 Th2 population against
LSODA at 1e-13 (tests/golden/th2_uniform_full.npz), by integrator and tolerance."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
f = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_uniform_full.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
rows = [int(a) for a in sys.argv[1:]] or [902, 903, 191, 183, 258, 53, 0, 1]
tight = np.concatenate([f["tight_ctrl_end"], f["tight_knock_end"]], axis=1)
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
print("rows", rows, "NO_IF", os.environ.get("LASSIM_NO_IF"))
g128 = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_uniform.npz")))
well = np.abs(g128["ref_cost_main"] - g128["tight_cost_main"]) / g128["tight_cost_main"] <= 1e-7
for method in ("dop853", "dp45"):
    eng.set_method(method)
    for guard in (2.0, 0.9, 0.5, 0.25):
        for tol in (1.49012e-8, 1e-9):
            eng.set_jump_guard(guard)
            eng.set_tolerances(tol, tol)
            r = eng.eval(X, details=True)
            ms = eng.last_kernel_ms()
            E = eng.last_end_states(1024)[:, 1:]
            st = eng.last_traj_steps(1024)
            with np.errstate(all="ignore"):
                e = np.abs(E - tight) / np.maximum(np.abs(tight), p["y0"][None, None, :])
            em = e.max(axis=(1, 2))
            cm = np.abs(r["cost_main"] - f["tight_cost_main"]) / f["tight_cost_main"]
            c128 = (np.abs(r["cost_main"][:128] - g128["ref_cost_main"]) / g128["ref_cost_main"])[well].max()
            print("%-7s guard %.2f tol %.0e  ms %.1f  end states: rows over 1e-5 %3d max %.1e  attempts kd %d series %d "
                  "| cost_main vs tight: over 1e-6 %d max %.1e; golden 128 vs ref %.1e | selected:" % (
                      method, guard, tol, ms, (em > 1e-5).sum(), em.max(), st[:, 1:].sum(), st[:, 0].sum(),
                      (cm > 1e-6).sum(), cm.max(), c128),
                  " ".join("%.1e" % em[r_] for r_ in rows), flush=True)

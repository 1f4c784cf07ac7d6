import sys, time, numpy as np
import os; REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
e = Engine(0)
e.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
a = e.eval(np.tile(X, (32, 1)))      # above 16,384 candidates: standard kernel only, as below
t0 = time.time(); b = e.eval(np.tile(X, (128, 1)), details=True); dt = time.time() - t0
print("N=131072: %.2f s -> %.0f evals/s, status ok %s, equal to N=32768 run: %s, periodic: %s"
      % (dt, 131072 / dt, (b["status"] == 0).all(), np.array_equal(a[:1024], b["cost"][:1024], equal_nan=True),
         np.array_equal(b["cost"][:1024], b["cost"][-1024:], equal_nan=True)))

"""DEVELOPMENT TOOL: throughput of the Th2 evaluation with the machine full (16,384 candidates)
for a library variant (LASSIM_LIB=liblassim_xxx.so in lassim_b200/)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import _capi
if os.environ.get("LASSIM_LIB"):
    _capi._LIB_PATH = os.path.join(REPO, "lassim_b200", os.environ["LASSIM_LIB"])
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
for method in ("dop853", "dp45"):
    eng.set_method(method)
    for reps, mode in ((16, "auto"), (16, "off"), (1, "auto")):
        eng.set_speculation(mode)
        Xs = np.tile(X, (reps, 1))
        eng.eval(Xs); ms = []
        for _ in range(3):
            r = eng.eval(Xs, details=True); ms.append(eng.last_kernel_ms())
        att = int(r["counters"][:, :2].sum())
        print(os.environ.get("LASSIM_LIB", "default"), method, mode, "N=%d" % len(Xs), "ms %.2f" % min(ms),
              "evals/s %.0f" % (len(Xs) / min(ms) * 1e3), "attempts %d" % att, flush=True)

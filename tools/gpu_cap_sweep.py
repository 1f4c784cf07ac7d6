"""DEVELOPMENT TOOL: hand-over cap on populations of 1024 (the bench's seed-0 population and the
archipelago's seeds): evaluation time, time series handed over, attempts of the longest."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import _capi
if os.environ.get("LASSIM_LIB"):
    _capi._LIB_PATH = os.path.join(REPO, "lassim_b200", os.environ["LASSIM_LIB"])
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
caps = [int(a) for a in sys.argv[1:]] or [64, 128, 256, 384, 512, 768, 1024]
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
for seed in (0, 1000, 1001, 1005):
    X = np.random.default_rng(seed).uniform(p["lb"], p["ub"], size=(1024, 87))
    line = "seed %4d" % seed
    for cap in caps:
        eng.set_series_cap(cap)
        eng.eval(X); ms = []
        for _ in range(5):
            eng.eval(X); ms.append(eng.last_kernel_ms())
        st = eng.last_traj_steps(1024)[:, 0]
        line += "  cap %4d: %6.2f ms (%3d over, capped %4.2f, sp %5.2f)" % (cap, np.median(ms), int((st > cap).sum()), eng.last_capped_phase_ms(), eng.last_stage_parallel_ms())
    print(line, flush=True)

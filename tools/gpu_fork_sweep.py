"""DEVELOPMENT TOOL: evaluation time by batch size for the Th2 problem (uniform populations,
seed 0 and 1000), for A/B runs of how the knock-down kernel is started (environment)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
for seed in (0, 1000):
    Xall = np.random.default_rng(seed).uniform(p["lb"], p["ub"], size=(16384, 87))
    line = "seed %4d" % seed
    for N in (8, 64, 256, 512, 1024, 2048, 4096, 8192, 16384):
        X = Xall[:N]
        eng.eval(X); ms = []
        for _ in range(3):
            eng.eval(X); ms.append(eng.last_kernel_ms())
        line += "  N=%5d: %6.2f" % (N, np.median(ms))
    print(line, flush=True)

"""DEVELOPMENT TOOL: hand-over cap against batch size (evaluation time, Th2 populations)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from workloads import uniform_population
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
caps = [int(a) for a in sys.argv[1:]] or [250, 500, 1000, 2000, 4000, 8000]
for N in (64, 256, 1024, 4096, 8192, 16384, 65536):
    X = uniform_population(p["lb"], p["ub"], min(N, 16384), seed=0)
    if N > 16384: X = np.tile(X, (N // 16384, 1))
    line = f"N={N:6d}"
    for cap in ["off"] + caps:
        if cap != "off": os.environ["LASSIM_SERIES_CAP"] = str(cap)
        eng = Engine(0)
        eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
        eng.set_speculation("off" if cap == "off" else "handover")
        eng.eval(X); ms = []
        for _ in range(3):
            eng.eval(X); ms.append(eng.last_kernel_ms())
        line += f"  {cap}: {min(ms):7.2f}"
        eng.close()
    print(line, flush=True)

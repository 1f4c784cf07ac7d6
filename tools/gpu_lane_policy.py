"""DEVELOPMENT TOOL: lanes against lockstep for archipelagos of 128-individual islands on one
GPU (where the automatic choice of lassim_de_set_lanes should switch)."""
import json, os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
def run(islands, np_, gens, lanes, seed=77):
    X0 = np.random.default_rng(seed).uniform(p["lb"], p["ub"], size=(islands, np_, 87))
    eng.de_init(X0, p["lb"], p["ub"], variant=3, f=0.85, cr=1.0, seed=seed)
    eng.de_set_lanes(lanes)
    eng.de_evolve(1); eng.de_get()
    t0 = time.perf_counter()
    eng.de_evolve(gens); eng.de_get()
    return (time.perf_counter() - t0) * 1e3
for islands, np_ in ((4, 128), (8, 128), (12, 128), (16, 128), (24, 128), (32, 128), (4, 512), (2, 1024), (16, 64), (32, 64)):
    line = "%2d x %4d:" % (islands, np_)
    for lanes in (1, 2, 4, 8, 16):
        if lanes > islands: continue
        line += "  lanes %2d %7.1f ms" % (lanes, run(islands, np_, 10, lanes))
    print(line, flush=True)

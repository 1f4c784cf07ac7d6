"""DEVELOPMENT TOOL: how much the asynchronous island lanes buy, and why.
Evolve rounds of Th2 islands on one GPU with 1 lane (lockstep) and with lanes of their own;
one island alone (the chain of its generations is what a lane cannot shorten); and a census
of the longest time series in the evolved populations."""
import json, os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
res = {}
def run(islands, np_, gens, lanes, seed=77):
    X0 = np.random.default_rng(seed).uniform(p["lb"], p["ub"], size=(islands, np_, 87))
    eng.de_init(X0, p["lb"], p["ub"], variant=3, f=0.85, cr=1.0, seed=seed)
    eng.de_set_lanes(lanes)
    eng.de_evolve(1); eng.de_get()
    t0 = time.perf_counter()
    eng.de_evolve(gens); X, F = eng.de_get()
    dt = time.perf_counter() - t0
    return dt * 1e3, X
for islands, np_, lanes in ((1, 128, 1), (8, 128, 1), (8, 128, 8), (8, 128, 4), (32, 128, 1), (32, 128, 16),
                            (64, 128, 16), (64, 128, 32), (128, 64, 32), (1, 1024, 1), (8, 1024, 8), (8, 1024, 1)):
    ms, X = run(islands, np_, 20, lanes)
    key = "%dx%d lanes=%d" % (islands, np_, lanes)
    res[key] = {"ms": ms, "evals_per_s": islands * np_ * 20 / (ms * 1e-3)}
    print(key, json.dumps(res[key]), flush=True)
# census: longest time series per island population after 21 generations, and fresh uniform ones
ms, X = run(8, 128, 20, 8)
census = []
for i in range(8):
    eng.eval(X[i]); k_ms = eng.last_kernel_ms()
    st = eng.last_traj_steps(128)
    census.append({"island": i, "ms": k_ms, "series_max": int(st[:, 0].max()), "series_over_768": int((st[:, 0] > 768).sum()),
                   "series_over_3000": int((st[:, 0] > 3000).sum())})
res["evolved islands (21 generations)"] = census
fresh = []
for s in range(8):
    Xs = np.random.default_rng(500 + s).uniform(p["lb"], p["ub"], size=(128, 87))
    eng.eval(Xs); k_ms = eng.last_kernel_ms()
    st = eng.last_traj_steps(128)
    fresh.append({"seed": 500 + s, "ms": k_ms, "series_max": int(st[:, 0].max()), "series_over_768": int((st[:, 0] > 768).sum()),
                  "series_over_3000": int((st[:, 0] > 3000).sum())})
res["fresh uniform populations of 128"] = fresh
print(json.dumps(census)); print(json.dumps(fresh))
json.dump(res, open(os.path.join(REPO, "gpurun_out", "async_probe.json"), "w"), indent=1)

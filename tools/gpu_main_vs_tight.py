"""DEVELOPMENT TOOL: time-series cost of the time-series modes against the same candidates
integrated at rtol = atol = 1e-12, on several random Th2 populations of 1024."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200._capi import Engine
from workloads import uniform_population
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 1.0)
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    X = uniform_population(p["lb"], p["ub"], 1024, seed=seed)
    eng.set_tolerances(1e-12, 1e-12); eng.set_speculation("off")
    t = eng.eval(X, details=True)
    Et = eng.last_end_states(1024)[:, 1:]
    eng.set_tolerances()
    eng.set_speculation("auto")
    eng.eval(X)
    E = eng.last_end_states(1024)[:, 1:]
    with np.errstate(all="ignore"):
        ee = (np.abs(E - Et) / np.maximum(np.abs(Et), p["y0"][None, None, :])).max(axis=(1, 2))
    line = f"seed {seed}: knock-down end states vs the same at 1e-12: rows > 1e-5: {int((ee > 1e-5).sum())} max {np.nanmax(ee):.1e} |"
    for mode in ("off", "handover", "stages3"):
        eng.set_speculation(mode)
        a = eng.eval(X, details=True)
        d = np.abs(a["cost_main"] - t["cost_main"]) / np.abs(t["cost_main"])
        line += f"  {mode}: status!=0 {int((a['status'] != 0).sum())} max {np.nanmax(d):.1e} (row {int(np.nanargmax(d))}) >1e-5: {int((d > 1e-5).sum())} p99 {np.nanpercentile(d, 99):.1e} ms {eng.last_kernel_ms():.1f}"
    print(line, flush=True)

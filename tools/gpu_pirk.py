"""DEVELOPMENT TOOL: the three time-series kernels (standard / speculative /
stage-parallel) on the Th2 golden populations and the bench population:
time-series cost against the converged answer, time, step counts, and the two
longest time series alone."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200._capi import Engine
from workloads import uniform_population

G = os.path.join(REPO, "tests", "golden")
p = dict(np.load(os.path.join(G, "th2_problem.npz")))
eng = Engine(0)
MODES = ("off", "on", "stages", "stages2", "stages3", "handover")
for name in ("th2_uniform", "th2_late"):
    g = np.load(os.path.join(G, name + ".npz"))
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], float(g["pert_factor"]))
    for mode in MODES:
        eng.set_speculation(mode)
        out = eng.eval(g["X"], details=True)
        ref, tight = g["ref_cost_main"], g["tight_cost_main"]
        well = np.abs(ref - tight) / np.abs(tight) <= 1e-7
        r_ref = np.abs(out["cost_main"] - ref) / np.abs(ref)
        r_t = np.abs(out["cost_main"] - tight) / np.abs(tight)
        print(f"{name} {mode:7s} cost_main vs reference (well-conditioned) max {r_ref[well].max():.2e} row {np.argmax(np.where(well, r_ref, 0))}"
              f" | vs converged max {r_t.max():.2e} median {np.median(r_t):.1e} | status ok {(out['status'] == 0).all()}")

X = uniform_population(p["lb"], p["ub"], 1024, seed=0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 1.0)
costs = {}
for mode in MODES:
    eng.set_speculation(mode)
    costs[mode] = eng.eval(X)
    ms = []
    for _ in range(3):
        eng.eval(X); ms.append(eng.last_kernel_ms())
    steps = eng.last_traj_steps(1024)
    print(f"N=1024 {mode:7s} ms {min(ms):.2f}  series attempts max {steps[:,0].max()} (row {steps[:,0].argmax()}) sum {steps[:,0].sum()} all {steps.sum()}")
d = np.abs(costs["stages"] - costs["off"]) / np.abs(costs["off"])
print("N=1024 stages vs off: max rel %.2e (row %d), 99th pct %.1e" % (np.nanmax(d), np.nanargmax(d), np.nanpercentile(d, 99)))

eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
for row in (820, 415, 0):
    for mode in MODES:
        eng.set_speculation(mode)
        eng.trajectories(X[row]); Y = eng.trajectories(X[row])
        ms = eng.last_kernel_ms(); st = int(eng.last_traj_steps(1)[0, 0])
        print(f"lone row {row} {mode:7s} attempts {st:6d} ms {ms:7.3f} us/attempt {1e3*ms/st:.3f}")

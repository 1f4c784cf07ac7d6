"""DEVELOPMENT TOOL: critical-path latency (ns per step attempt of one lone
group) and bulk time of the Th2 population on a B200."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200._capi import Engine
gold = os.path.join(REPO, "tests", "golden")
p = dict(np.load(os.path.join(gold, "th2_problem.npz")))
g = dict(np.load(os.path.join(gold, "th2_uniform.npz")))
eng = Engine(0)
eng.set_method(os.environ.get('LASSIM_METHOD', 'dop853'))
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
r = eng.eval(X, details=True)
steps = eng.last_traj_steps(1024)
ms = [0, 0, 0]
for i in range(3):
    eng.eval(X); ms[i] = eng.last_kernel_ms()
worst = int(np.argmax(steps[:, 0]))
print("N=1024 full: kernel ms", ["%.2f" % m for m in ms], "attempts", int(steps.sum()), "longest", int(steps.max()))
for rep in range(2):
    Y = eng.trajectories(X[worst])
    k = eng.last_kernel_ms()
    s1 = eng.last_traj_steps(1)[0, 0]
    print("lone group: %d steps in %.2f ms -> %.1f ns/step" % (s1, k, 1e6 * k / s1))
# 2 lone groups in one warp (same candidate twice)
Y = eng.trajectories(np.stack([X[worst], X[worst]]))
k = eng.last_kernel_ms(); print("two groups/one warp: %.2f ms" % k)
# a full machine of identical long trajectories: throughput-bound regime
Xl = np.repeat(X[worst][None], 148 * 3 * 8 * 2, axis=0)
Y = eng.trajectories(Xl)
k = eng.last_kernel_ms(); tot = eng.last_traj_steps(len(Xl))[:, 0].sum()
print("saturated (%d identical long trajectories): %.2f ms, %.3f ns per step-attempt machine-wide, %.2f TFLOP/s algorithmic"
      % (len(Xl), k, 1e6 * k / tot, tot * 1836 / (k * 1e-3) / 1e12))
gcm = g["ref_cost_main"]; well = np.abs(gcm - g["tight_cost_main"]) / g["tight_cost_main"] <= 1e-7
print("parity gated max rel", (np.abs(r["cost_main"][:128] - gcm) / gcm)[well].max())

"""DEVELOPMENT TOOL: first look at the CUDA path on a B200 - parity against
the golden vectors, kernel time, step statistics.  Writes gpurun_out/first_look.txt"""
import os
import sys
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200._capi import Engine  # noqa: E402

gold = os.path.join(REPO, "tests", "golden")
p = dict(np.load(os.path.join(gold, "th2_problem.npz")))
out = []


def log(*a):
    s = " ".join(str(x) for x in a)
    print(s, flush=True)
    out.append(s)


eng = Engine(0)
eng.set_method(os.environ.get('LASSIM_METHOD', 'dop853'))
log("fp64 probe TFLOP/s:", eng.fp64_peak_tflops(200000))
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
log("dim", eng.dim, "K", eng.trajectories_per_candidate)
for name in ("th2_uniform", "th2_late"):
    g = dict(np.load(os.path.join(gold, name + ".npz")))
    X = g["X"]
    t0 = time.time()
    r = eng.eval(X, details=True)
    dt = time.time() - t0
    ms = eng.last_kernel_ms()
    ref, tight = g["ref_cost_main"], g["tight_cost_main"]
    rel_ref = np.abs(r["cost_main"] - ref) / np.abs(ref)
    rel_tight = np.abs(r["cost_main"] - tight) / np.abs(tight)
    ref_tight = np.abs(ref - tight) / np.abs(tight)
    well = ref_tight <= 1e-7
    log(name, "N", len(X), "wall %.3fs kernel %.2f ms" % (dt, ms), "status!=0:", int((r["status"] != 0).sum()))
    log("  cost_main rel vs ref: med %.2e max %.2e ; vs tight med %.2e max %.2e ; gated n=%d max %.2e"
        % (np.median(rel_ref), rel_ref.max(), np.median(rel_tight), rel_tight.max(), well.sum(), rel_ref[well].max()))
    relp = np.abs(r["pert_cost"] - g["tight_pert_cost"]) / np.abs(g["tight_pert_cost"])
    log("  pert rel vs tight: med %.2e max %.2e ; ref-vs-tight med %.2e"
        % (np.nanmedian(relp), np.nanmax(relp),
           np.median(np.abs(g["ref_pert_cost"] - g["tight_pert_cost"]) / g["tight_pert_cost"])))
    steps = eng.last_traj_steps(len(X))
    log("  steps main med %d max %d ; per-candidate total med %d max %d ; rej frac %.3f"
        % (np.median(steps[:, 0]), steps[:, 0].max(), np.median(steps.sum(1)), steps.sum(1).max(),
           r["counters"][:, 1].sum() / r["counters"][:, :2].sum()))
    Y = eng.trajectories(X)
    sj = np.abs(g["tight_Y"]).max(axis=1, keepdims=True)
    dY = np.abs(Y - g["tight_Y"]) / np.maximum(np.abs(g["tight_Y"]), sj)
    log("  traj mixed rel err vs tight: max %.2e" % dY.max())

# full-size population
lb, ub = p["lb"], p["ub"]
X = np.random.default_rng(0).uniform(lb, ub, size=(1024, lb.size))
for rep in range(3):
    t0 = time.time()
    r = eng.eval(X, details=True)
    dt = time.time() - t0
    log("N=1024 rep", rep, "wall %.3f s kernel %.2f ms -> %.0f evals/s" % (dt, eng.last_kernel_ms(), 1024 / dt))
steps = eng.last_traj_steps(1024)
tot = r["counters"][:, :2].sum()
log("  attempts total %d ; main med %d max %d ; longest traj %d ; status!=0 %d"
    % (tot, np.median(steps[:, 0]), steps[:, 0].max(), steps.max(), int((r["status"] != 0).sum())))
flops = tot * (12 * 63 + 90 * 12)
log("  algorithmic FP64 flops %.3e -> %.3f TFLOP/s" % (flops, flops / (eng.last_kernel_ms() * 1e-3) / 1e12))
os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
tag = os.environ.get("LASSIM_GROUP", "default")
open(os.path.join(REPO, "gpurun_out", "first_look_%s.txt" % tag), "w").write("\n".join(out) + "\n")

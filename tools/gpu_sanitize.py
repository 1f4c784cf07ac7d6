"""DEVELOPMENT TOOL: a tiny evaluation of every kernel family (written for compute-sanitizer,
which this pool has since closed; run plainly it shows that all kernels agree on small cases)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200 import Engine
from workloads import synthetic_problem
th2 = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
g = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_late.npz")))
for method in ("dop853", "dp45"):
    eng = Engine(0); eng.set_method(method)
    eng.set_tolerances(1e-5, 1e-5)          # few steps: the sanitizer is slow
    eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], th2["pert"], 0.5)
    r = eng.eval(g["X"][:5], details=True); print(method, "th2", r["cost"][:2], r["status"])
    eng.trajectories(g["X"][:3])
    eng.set_peripheral(g["X"][0], np.random.default_rng(0).uniform(0.1, 1, (4, 6)), 0.2, np.full(4, 0.5))
    print(method, "periph", eng.eval_peripheral(np.arange(4), np.random.default_rng(1).uniform(0, 1, (4, 14))))
    for n in (50, 200):
        p = synthetic_problem(n, seed=n)
        eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
        X = np.random.default_rng(2).uniform(p["lb"], p["ub"], (3, p["lb"].size)) * 0.1
        print(method, n, eng.eval(X))
    eng.close()

# ---- kernel families of round 2: every time-series kernel mode (hand-over forced by a cap of
# 8 attempts), the wide stage-parallel kernel (n = 50), core table + peripheral march (and the
# joint kernel), device DE with island lanes, the entry point that allocates nothing
import torch
eng = Engine(0)
eng.set_tolerances(1e-5, 1e-5)
eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], th2["pert"], 0.5)
for mode in ("off", "on", "stages", "stages2", "stages3", "handover"):
    eng.set_speculation(mode)
    eng.set_series_cap(8 if mode == "handover" else 0)
    r = eng.eval(g["X"][:6], details=True)
    print("mode", mode, r["cost"][:2], r["status"], "stage-parallel attempts", eng.last_stage_parallel_attempts())
eng.set_speculation("auto"); eng.set_series_cap(0)
rng = np.random.default_rng(3)
pd_, ps_ = rng.uniform(0.1, 1, (5, 6)), np.full(5, 0.5)
Xp = rng.uniform(0, 1, (7, 14))
for joint in (False, True):
    eng.set_peripheral_joint(joint)
    eng.set_peripheral(g["X"][0], pd_, 0.2, ps_)
    print("periph joint", joint, eng.eval_peripheral(np.arange(7) % 5, Xp)[:3])
    eng.peripheral_trajectories(np.arange(3), Xp[:3])
eng.set_peripheral_joint(False)
lb, ub = th2["lb"], th2["ub"]
X0 = rng.uniform(lb, ub, (4, 8, lb.size)) * 0.2
for lanes in (1, 2):
    eng.de_init(X0, lb, ub, variant=3, f=0.85, cr=1.0, seed=1)
    eng.de_set_lanes(lanes)
    eng.de_evolve(2)
    tb = eng.de_best()
    eng.de_migrate_ring(tb)
    eng.de_evolve(1)
    print("de lanes", lanes, eng.de_get()[1].min(axis=1))
dX = torch.from_numpy(g["X"][:6]).cuda()
dC = torch.empty(6, dtype=torch.float64, device="cuda")
eng.eval_device_noalloc(6, dX.data_ptr(), dC.data_ptr())
torch.cuda.synchronize()
print("noalloc", dC.cpu().numpy()[:2])
eng.close()
eng = Engine(0); eng.set_tolerances(1e-5, 1e-5)
p = synthetic_problem(50, seed=50)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
X = np.random.default_rng(2).uniform(p["lb"], p["ub"], (3, p["lb"].size)) * 0.1
for mode in ("stages", "handover"):
    eng.set_speculation(mode); eng.set_series_cap(8 if mode == "handover" else 0)
    print("wide", mode, eng.eval(X), eng.last_stage_parallel_attempts())
eng.close()

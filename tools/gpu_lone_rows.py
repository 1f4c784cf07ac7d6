"""DEVELOPMENT TOOL: the slowest time series of the bench population evaluated (almost) alone:
capped phase and stage-parallel kernel without a population around them."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
eng.eval(X)
st = eng.last_traj_steps(1024)[:, 0]
order = np.argsort(-st.astype(np.int64))
print("longest rows", order[:6].tolist(), st[order[:6]].tolist())
for rows in (order[:1], order[:2], order[:8], order[:64]):
    eng.eval(X[rows]); ms = []
    for _ in range(3):
        eng.eval(X[rows]); ms.append(eng.last_kernel_ms())
    print("%3d rows: %.2f ms (capped %.2f, stage-parallel %.2f, attempts of the stage-parallel kernel %d)"
          % (len(rows), np.median(ms), eng.last_capped_phase_ms(), eng.last_stage_parallel_ms(),
             eng.last_stage_parallel_attempts()), flush=True)

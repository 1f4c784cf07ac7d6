"""DEVELOPMENT TOOL: per-attempt latency of the time-series kernels on BASELINE config 3's
problem (50 genes): the slowest candidates alone, and the whole population, by mode and cap."""
import json, os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from workloads import synthetic_problem, uniform_population
from lassim_b200 import Engine
p = synthetic_problem(50, seed=50)
X = uniform_population(p["lb"], p["ub"], 1024, seed=50)
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
eng.set_speculation("off")
eng.eval(X)
st = eng.last_traj_steps(1024)[:, 0]
order = np.argsort(-st)
print("series attempts (standard): max", st.max(), "median", int(np.median(st)), "over 768:", int((st > 768).sum()),
      "over 3000:", int((st > 3000).sum()), "sum", int(st.sum()))
for rows in ([order[0]], [order[1]], list(order[:8]), list(order[:148]), list(range(1024))):
    Xs = X[rows]
    line = "N=%4d" % len(rows)
    for mode, cap in (("off", 0), ("stages", 0), ("handover", 0), ("handover", 200), ("handover", 3000)):
        eng.set_speculation(mode); eng.set_series_cap(cap)
        eng.eval(Xs); ms = []
        for _ in range(3):
            eng.eval(Xs); ms.append(eng.last_kernel_ms())
        s = eng.last_traj_steps(len(rows))[:, 0]
        extra = ""
        if mode != "off":
            try:
                extra = " sp %.2f ms %d att" % (eng.last_stage_parallel_ms(), eng.last_stage_parallel_attempts())
            except Exception:
                pass
        line += " | %s/%d: %.2f ms max att %d%s" % (mode, cap, min(ms), s.max(), extra)
    print(line, flush=True)

# the whole evaluation (52 solves per candidate), by cap
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
for mode, cap in (("off", 0), ("handover", 768), ("handover", 1500), ("handover", 3000), ("handover", 6000)):
    eng.set_speculation(mode); eng.set_series_cap(cap)
    eng.eval(X); ms = []
    for _ in range(3):
        eng.eval(X); ms.append(eng.last_kernel_ms())
    extra = ""
    if mode != "off":
        extra = " sp %.2f ms %d att" % (eng.last_stage_parallel_ms(), eng.last_stage_parallel_attempts())
    print("full evaluation N=1024 %s/%d: %.2f ms%s" % (mode, cap, min(ms), extra), flush=True)

"""DEVELOPMENT STUDY (CPU): several step sizes attempted side by side with the
stage-parallel Gauss pair (the WARPS = 2, 3 variants of pirk_kernel.cuh): sequential
attempts on the two limit-cycle time series of the bench population.  Not part of the
product or the tests.

  python tools/study/pirk_sizes_study.py            the ratio sets quoted in pirk_kernel.cuh
  python tools/study/pirk_sizes_study.py 1.2 1 0.7  one set
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import pirk_study as ps  # noqa: E402
from workloads import uniform_population  # noqa: E402

times, y0, TOL = ps.times, ps.y0, ps.TOL


def attempt(f, sig, A, b, c, polys, y, prev, hs, s, m, order, swing_limit):
    if prev is None:
        F = np.tile(f(y), (s, 1))
    else:
        yp, hp, Fp = prev
        th = 1 + c * hs / hp
        Pm = np.array([[polys[j](th[i]) - polys[j](1.0) for j in range(s)] for i in range(s)])
        F = f(y + hp * (Pm @ Fp))
    Fo = F
    for _ in range(m):
        Fo = F; Y = y + hs * (A @ F); F = f(Y)
    yn = y + hs * (b @ F); yl = y + hs * (b @ Fo)
    sc = TOL + TOL * np.maximum(np.abs(y), np.abs(yl))
    err = np.max(np.abs(yn - yl) / sc)
    sg = sig(Y)
    swing = np.max(sg.max(axis=0) - sg.min(axis=0))
    if swing > swing_limit:
        err = max(err, (swing / swing_limit) ** order)
    return yn, F, (err if np.isfinite(err) else 1e30)


def run(x, ratios, s=5, m=5, swing_limit=0.5):
    """The controller of grn_series_pirk_kernel: the largest attempt that passes is taken; a
    larger one having failed forbids growth at the next step."""
    f, sig = ps.system(x)
    order = 2 * s
    A, b, c, polys = ps.gauss(s)
    t = times[0]; y = y0.copy()
    h = 0.01 * np.max(np.abs(y) + 1e-6) / max(np.max(np.abs(f(y))), 1e-12)
    seq = 0; prev = None; rejected = False; used = [0] * len(ratios)
    for tend in times[1:]:
        while t < tend:
            seq += 1
            res = []
            for r in ratios:
                hs = min(h * r, tend - t)
                res.append((hs,) + attempt(f, sig, A, b, c, polys, y, prev, hs, s, m, order, swing_limit))
            pick = next((i for i, v in enumerate(res) if v[3] <= 1), None)
            if pick is None:
                hs, _, _, err = res[-1]
                h = hs * min(1.0, max(0.2, 0.9 * err ** (-1.0 / order))); rejected = True
            else:
                hs, yn, F, err = res[pick]; used[pick] += 1
                fac = 10.0 if err == 0 else min(10.0, max(0.2, 0.9 * err ** (-1.0 / order)))
                if rejected or pick > 0: fac = min(fac, 1.0)
                rejected = pick > 0
                prev = (y, hs, F); t += hs; y = yn; h = hs * fac
    return seq, used


if __name__ == "__main__":
    X = uniform_population(ps.g["lb"], ps.g["ub"], 1024, seed=0)
    sets = [tuple(float(a) for a in sys.argv[1:])] if len(sys.argv) > 1 else \
        [(1.0,), (1.0, 0.7), (1.1, 1.0, 0.7), (1.15, 1.0, 0.7), (1.2, 1.0, 0.7), (1.25, 1.0, 0.75),
         (1.0, 0.85, 0.7, 0.5)]
    for ratios in sets:
        for row in (415, 820):
            seq, used = run(X[row], ratios)
            print("row", row, "ratios", ratios, "sequential attempts", seq, "taken per size", used, flush=True)

"""DEVELOPMENT STUDY (CPU): a sharper error estimate for the stage-parallel Gauss pair.
The kernel takes the last correction |y(M) - y(M-1)| as the error of y(M); that is the error of
y(M-1).  With the contraction rho = |y(M) - y(M-1)| / |y(M-1) - y(M-2)| the error of y(M) is about
rho / (1 - rho) times the last correction.  Attempts and accuracy with either estimate on the
golden sets and the two limit cycles.  Not part of the product or the tests.

  python tools/study/pirk_estimator_study.py [floor]    floor = smallest factor applied (default 0.1)
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import pirk_study as ps  # noqa: E402
from workloads import uniform_population  # noqa: E402

times, y0, TOL = ps.times, ps.y0, ps.TOL


def pirk(x, sharp, floor, s=5, m=5, swing_limit=0.5, facmax=10.0, safety=0.9):
    f, sig = ps.system(x)
    order = 2 * s
    A, b, c, polys = ps.gauss(s)
    t = times[0]; y = y0.copy()
    f0 = f(y)
    h = 0.01 * np.max(np.abs(y) + 1e-6) / max(np.max(np.abs(f0)), 1e-12)
    acc = rej = 0
    prev = None; ends = []; rejected = False
    for tend in times[1:]:
        while t < tend:
            hs = min(h, tend - t)
            if prev is None:
                F = np.tile(f(y), (s, 1))
            else:
                yp, hp, Fp = prev
                th = 1 + c * hs / hp
                Pm = np.array([[polys[j](th[i]) - polys[j](1.0) for j in range(s)] for i in range(s)])
                F = f(y + hp * (Pm @ Fp))
            Fo = Foo = F
            for k in range(m):
                Foo = Fo; Fo = F; Y = y + hs * (A @ F); F = f(Y)
            yn = y + hs * (b @ F); yl = y + hs * (b @ Fo); yll = y + hs * (b @ Foo)
            sc = TOL + TOL * np.maximum(np.abs(y), np.abs(yl))
            d1 = np.max(np.abs(yn - yl) / sc); d0 = np.max(np.abs(yl - yll) / sc)
            err = d1
            if sharp and d0 > 0 and np.isfinite(d0):
                rho = d1 / d0
                if rho < 0.9:
                    err = d1 * min(1.0, max(floor, rho / (1 - rho)))
            sg = sig(Y)
            swing = np.max(sg.max(axis=0) - sg.min(axis=0))
            if swing > swing_limit:
                err = max(err, (swing / swing_limit) ** order)
            if not np.isfinite(err): err = 1e30
            fac = facmax if err == 0 else min(facmax, max(0.2, safety * err ** (-1.0 / order)))
            if err <= 1:
                prev = (y, hs, F); t += hs; y = yn; acc += 1
                if rejected: fac = min(fac, 1.0)
                rejected = False; h = hs * fac
            else:
                rej += 1; h = hs * min(1.0, fac); rejected = True
        ends.append(y.copy())
    return np.array(ends), acc, rej


if __name__ == "__main__":
    floor = float(sys.argv[1]) if len(sys.argv) > 1 else 0.1
    G = ps.G
    for name in ("th2_late", "th2_uniform"):
        d = np.load(os.path.join(G, name + ".npz"))
        for sharp in (False, True):
            errs = []; att = 0
            for i in range(0, d["X"].shape[0], 2):
                E, acc, rej = pirk(d["X"][i], sharp, floor)
                ct = d["tight_cost_main"][i]
                conv = abs(d["ref_cost_main"][i] - ct) / ct <= 1e-7
                errs.append((abs(ps.cost(E) - ct) / ct, conv)); att += acc + rej
            e_conv = [e for e, c in errs if c]
            print(name, "sharp" if sharp else "plain", "attempts", att, "max err (converged rows) %.1e" % max(e_conv),
                  "median %.1e" % np.median(e_conv), "max err (all) %.1e" % max(e for e, _ in errs), flush=True)
    X = uniform_population(ps.g["lb"], ps.g["ub"], 1024, seed=0)
    for sharp in (False, True):
        E, acc, rej = pirk(X[415], sharp, floor)
        print("row 415", "sharp" if sharp else "plain", "attempts", acc + rej, "rejected %.2f" % (rej / (acc + rej)), flush=True)

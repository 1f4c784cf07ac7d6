"""DEVELOPMENT STUDY (CPU): the stage-parallel Gauss pair of pirk_kernel.cuh
(continued-collocation predictor, M corrections, error estimate = last correction,
logistic-swing limit) against Dormand-Prince 8(5,3) on the Th2 golden populations:
sequential right-hand sides ("depth") and error of the time-series cost against the
converged answer.  Not part of the product or the tests.

  python tools/study/pirk_study.py golden [swing]     all golden rows, cost error
  python tools/study/pirk_study.py depth [swing]      the two limit cycles of the bench population
"""
import os
import sys

import numpy as np
from scipy.integrate import solve_ivp

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from workloads import uniform_population  # noqa: E402

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests", "golden")
g = np.load(os.path.join(G, "th2_problem.npz"))
n = g["y0"].size
mask = g["mask"].reshape(n, n).astype(bool)
y0 = g["y0"].astype(float)
times = g["times"].astype(float)
TOL = 1.49012e-8


def system(x):
    lam, vm = x[:n], x[n:2 * n]
    W = np.zeros((n, n)); W[mask] = x[2 * n:]
    WT = W.T.copy()

    def sig(Y):
        with np.errstate(over="ignore"):
            return 1 / (1 + np.exp(-(Y @ WT)))

    def f(Y):
        return -lam * Y + vm * sig(Y)
    return f, sig


def gauss(s):
    c, w = np.polynomial.legendre.leggauss(s)
    c = (c + 1) / 2; b = w / 2
    A = np.zeros((s, s)); polys = []
    for j in range(s):
        others = [c[k] for k in range(s) if k != j]
        p = np.poly1d(np.poly(others)) / np.prod([c[j] - o for o in others])
        P = p.integ(); polys.append(P)
        A[:, j] = [P(ci) for ci in c]
    return A, b, c, polys


def cost(Y):
    Yf = np.vstack([y0[None, :], Y])
    return np.sum((g["data"] - Yf / Yf.max(axis=0)) ** 2 / g["sigma"] ** 2)


def pirk(x, s=5, m=5, swing_limit=None, rtol=TOL, atol=TOL, facmax=10.0, safety=0.9, pi_beta=0.0):
    f, sig = system(x)
    order = 2 * s
    A, b, c, polys = gauss(s)
    t = times[0]; y = y0.copy()
    f0 = f(y)
    h = 0.01 * np.max(np.abs(y) + 1e-6) / max(np.max(np.abs(f0)), 1e-12)
    acc = rej = depth = 0
    prev = None; ends = []; rejected = False; err_old = 1.0
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
            Fo = F
            for k in range(m):
                Fo = F; Y = y + hs * (A @ F); F = f(Y)
            depth += m + 1
            yn = y + hs * (b @ F)
            yl = y + hs * (b @ Fo)
            sc = atol + rtol * np.maximum(np.abs(y), np.abs(yl))
            err = np.max(np.abs(yn - yl) / sc)
            if swing_limit is not None:
                sg = sig(Y)
                swing = np.max(sg.max(axis=0) - sg.min(axis=0))
                if swing > swing_limit:
                    err = max(err, (swing / swing_limit) ** order)
            if not np.isfinite(err): err = 1e30
            fac = facmax if err == 0 else min(facmax, max(0.2, safety * err ** (-(1.0 / order - 0.75 * pi_beta)) * err_old ** (-pi_beta)))
            if err <= 1:
                prev = (y, hs, F); t += hs; y = yn; acc += 1
                if rejected: fac = min(fac, 1.0)
                rejected = False; h = hs * fac; err_old = max(err, 1e-4)
            else:
                rej += 1; h = hs * min(1.0, fac); rejected = True
        ends.append(y.copy())
    return np.array(ends), acc, rej, depth


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "golden"
    lim = float(sys.argv[2]) if len(sys.argv) > 2 else None
    if what == "golden":
        for name in ("th2_late", "th2_uniform"):
            d = np.load(os.path.join(G, name + ".npz"))
            errs = []
            for i in range(d["X"].shape[0]):
                E, acc, rej, depth = pirk(d["X"][i], swing_limit=lim)
                ct = d["tight_cost_main"][i]
                errs.append((abs(cost(E) - ct) / ct, i, abs(d["ref_cost_main"][i] - ct) / ct))
            errs.sort(reverse=True)
            print(name, "swing", lim, "worst:", ["row %d %.1e (ref %.0e)" % (i, e, r) for e, i, r in errs[:6]], flush=True)
    else:
        X = uniform_population(g["lb"], g["ub"], 1024, seed=0)
        for row in (415, 820):
            f, _ = system(X[row])
            ref = solve_ivp(lambda t, y: f(y), (times[0], times[-1]), y0, method="DOP853", rtol=1e-13, atol=1e-13, t_eval=times[1:])
            d8 = solve_ivp(lambda t, y: f(y), (times[0], times[-1]), y0, method="DOP853", rtol=TOL, atol=TOL, t_eval=times[1:])
            E, acc, rej, depth = pirk(X[row], swing_limit=lim)
            ct = cost(ref.y.T)
            print(f"row {row}: dop853 depth {d8.nfev} cost err {abs(cost(d8.y.T)-ct)/ct:.1e} | pirk swing {lim}: attempts {acc+rej} "
                  f"rej {rej/(acc+rej):.2f} depth {depth} cost err {abs(cost(E)-ct)/ct:.1e}", flush=True)

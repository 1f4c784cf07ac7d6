"""DEVELOPMENT TOOL: run tools/dp45_model.c (scalar model of the device step
controller) over a golden population and report step counts and parity against
the reference / tight costs.  Not part of the product or the test suite."""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
lib = ctypes.CDLL(os.path.join(HERE, "_dp45_model.so"))
dp = ctypes.POINTER(ctypes.c_double)
lib.dp45_integrate.restype = ctypes.c_int
lib.dp45_integrate.argtypes = [ctypes.c_int, dp, dp, dp, dp, ctypes.c_int, dp,
                               ctypes.c_double, ctypes.c_double, ctypes.c_long,
                               ctypes.c_double, dp, ctypes.POINTER(ctypes.c_long)]


def P(a):
    return a.ctypes.data_as(dp)


def integrate(lam, vmax, W, y0, ts, rtol, atol, beta=0.0, max_steps=10**8):
    n = y0.size
    lam = np.ascontiguousarray(lam); vmax = np.ascontiguousarray(vmax)
    W = np.ascontiguousarray(W); ts = np.ascontiguousarray(ts, dtype=np.float64)
    Y = np.empty((ts.size, n))
    cnt = (ctypes.c_long * 3)()
    st = lib.dp45_integrate(n, P(lam), P(vmax), P(W), P(y0), ts.size, P(ts),
                            rtol, atol, max_steps, beta, P(Y), cnt)
    return Y, st, list(cnt)


def evaluate(x, prob, rtol, atol, beta=0.0):
    n = prob["y0"].size
    mask = prob["mask"].reshape(n, n)
    lam = x[:n].copy(); vmax = x[n:2 * n].copy()
    W = np.zeros((n, n)); W[mask] = x[2 * n:]
    y0 = np.ascontiguousarray(prob["y0"])
    Y, st, c = integrate(lam, vmax, W, y0, prob["times"], rtol, atol, beta)
    with np.errstate(all="ignore"):
        cost = np.sum((prob["data"] - Y / Y.max(axis=0)) ** 2 / prob["sigma"] ** 2)
    steps = [c[0] + c[1]]
    rej = [c[1]]
    # perturbations
    pert = prob["pert"]
    Pm = pert[:, :n]
    tms = pert[:, n:]
    lamp = lam.copy()
    ctrl = {}
    sim = np.empty(n * n)
    for i in range(n):
        key = str(tms[i])
        if key not in ctrl:
            Yc, st2, c2 = integrate(lamp, vmax, W, y0, tms[i], rtol, atol, beta)
            ctrl[key] = Yc[-1]
            steps.append(c2[0] + c2[1]); rej.append(c2[1])
        lamp[i] *= Pm[i, i]
        Yk, st3, c3 = integrate(lamp, vmax, W, y0, tms[i], rtol, atol, beta)
        steps.append(c3[0] + c3[1]); rej.append(c3[1])
        with np.errstate(all="ignore"):
            sim[i * n:(i + 1) * n] = Yk[-1] / ctrl[key]
    sim -= 1
    sim[sim > 2] = 2
    p = Pm.flatten()
    with np.errstate(all="ignore"):
        m = np.max(np.abs(sim))
        r = sim / m
        pc = p @ p - (r @ p) ** 2 / (r @ r)
    return cost, pc, Y, steps, rej, min(np.min(np.abs(v)) for v in ctrl.values())


if __name__ == "__main__":
    gold = os.path.join(os.path.dirname(HERE), "tests", "golden")
    name = sys.argv[1] if len(sys.argv) > 1 else "th2_uniform"
    rtol = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-10
    beta = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
    nmax = int(sys.argv[4]) if len(sys.argv) > 4 else 10**9
    prob = dict(np.load(os.path.join(gold, "th2_problem.npz")))
    g = np.load(os.path.join(gold, name + ".npz"))
    X = g["X"][:nmax]
    res = [evaluate(x, prob, rtol, rtol, beta) for x in X]
    cost = np.array([r[0] for r in res]); pc = np.array([r[1] for r in res])
    ref = g["ref_cost_main"][:nmax]; tight = g["tight_cost_main"][:nmax]
    rel_ref = np.abs(cost - ref) / np.abs(ref)
    rel_tight = np.abs(cost - tight) / np.abs(tight)
    ref_tight = np.abs(ref - tight) / np.abs(tight)
    print("cost_main  rel vs ref  : med %.2e max %.2e  frac<=1e-6 %.3f" % (np.median(rel_ref), rel_ref.max(), np.mean(rel_ref <= 1e-6)))
    print("cost_main  rel vs tight: med %.2e max %.2e" % (np.median(rel_tight), rel_tight.max()))
    print("ref vs tight           : med %.2e max %.2e" % (np.median(ref_tight), ref_tight.max()))
    well = ref_tight <= 1e-7
    print("gated (ref~tight<=1e-7): n=%d  max rel vs ref %.2e" % (well.sum(), rel_ref[well].max()))
    main_steps = np.array([r[3][0] for r in res])
    pert_steps = np.array([sum(r[3][1:]) for r in res])
    maxpert = np.array([max(r[3][1:]) for r in res])
    rejm = np.array([r[4][0] for r in res])
    print("main steps: med %d mean %d max %d (rej frac %.3f)" % (np.median(main_steps), main_steps.mean(), main_steps.max(), rejm.sum() / main_steps.sum()))
    print("pert steps (14 traj): med %d mean %d max %d ; longest single pert traj: med %d max %d" % (np.median(pert_steps), pert_steps.mean(), pert_steps.max(), np.median(maxpert), maxpert.max()))
    print("lsoda nst med %d max %d" % (np.median(g["lsoda_nst"][:nmax]), g["lsoda_nst"][:nmax].max()))
    tp = g["tight_pert_cost"][:nmax]; rp = g["ref_pert_cost"][:nmax]
    cmin = np.array([r[5] for r in res])
    okc = cmin > 1e-6
    relp = np.abs(pc - tp) / np.abs(tp)
    print("pert: candidates with controls>1e-6: %d ; rel vs tight there: max %.2e ; overall med %.2e max %.2e" % (okc.sum(), relp[okc].max() if okc.any() else -1, np.median(relp), np.nanmax(relp)))
    print("pert ref vs tight overall: med %.2e max %.2e" % (np.median(np.abs(rp - tp) / tp), np.max(np.abs(rp - tp) / tp)))

"""
Seeded workloads named by BASELINE.json's configs.  Used by
the tests, the tools and the golden-vector generator; not part of the product
package.

  synthetic_problem(50, seed=50)   config 3: 50 genes, in-degree 5, nnz 250
  synthetic_problem(200, seed=200) config 4: 200 genes, in-degree 5, nnz 1000
  uniform_population(lb, ub, N, seed)  U[lb, ub] rows (config 2 uses seed 0)
  stiff_population(problem, N, seed)   config 4's lambda,vmax ~ U[10,20]

The "expression data" of a synthetic problem is the normalised trajectory of a
hidden ground-truth candidate, integrated once with SciPy's LSODA at 1e-12 when
the problem is generated (workload construction, not cost evaluation).
"""
import numpy as np

TH2_TIMES = np.array([0.0, 6.0, 24.0, 72.0, 144.0, 192.0])


def default_bounds_arrays(n, nnz):
    """lambda, vmax in [0, 20]; k in [-20, 20] (reference
    source/customs/core_functions.py:19-35)."""
    lb = np.concatenate([np.zeros(2 * n), np.full(nnz, -20.0)])
    ub = np.full(2 * n + nnz, 20.0)
    return lb, ub


def uniform_population(lb, ub, count, seed=0):
    rng = np.random.default_rng(seed)
    return rng.uniform(np.asarray(lb), np.asarray(ub),
                       size=(count, len(lb)))


def synthetic_problem(n, seed, in_degree=5, times=TH2_TIMES, sigma=0.2):
    rng = np.random.default_rng(seed)
    mask = np.zeros((n, n), dtype=bool)
    for i in range(n):
        mask[i, rng.choice(n, size=in_degree, replace=False)] = True
    nnz = int(mask.sum())
    y0 = rng.uniform(0.05, 1.5, size=n)
    # hidden ground truth: mild parameters so the series is not flat
    lam = rng.uniform(0.02, 0.5, size=n)
    vmax = rng.uniform(0.02, 1.0, size=n)
    k = rng.uniform(-3.0, 3.0, size=nnz)
    truth = np.concatenate([lam, vmax, k])
    W = np.zeros((n, n))
    W[mask] = k

    def rhs(y, t):
        with np.errstate(over="ignore"):
            return -lam * y + vmax / (1.0 + np.exp(-(W @ y)))

    from scipy.integrate import odeint
    Y = odeint(rhs, y0, np.asarray(times, dtype=np.float64),
               rtol=1e-12, atol=1e-12, mxstep=10 ** 7)
    data = Y / np.amax(Y, axis=0)
    pert = rng.uniform(-1.0, 2.0, size=(n, n + 2))
    pert[np.arange(n), np.arange(n)] = rng.uniform(-1.0, -0.05, size=n)
    pert[:, n] = 0.0
    pert[:, n + 1] = 24.0
    lb, ub = default_bounds_arrays(n, nnz)
    vector_map = np.where(mask.reshape(-1), -1.0, 0.0)
    return dict(data=data, sigma=np.full(n, float(sigma)),
                times=np.asarray(times, dtype=np.float64), y0=y0, pert=pert,
                vector_map=vector_map, mask=mask.reshape(-1).copy(),
                lb=lb, ub=ub, truth=truth, react_count=np.int64(nnz))


def stiff_population(problem, count, seed=200):
    """Config 4: lambda ~ U[10,20], vmax ~ U[10,20], k ~ U[-20,20]."""
    n = problem["y0"].size
    nnz = int(np.asarray(problem["mask"]).sum())
    rng = np.random.default_rng(seed)
    return np.concatenate([rng.uniform(10.0, 20.0, size=(count, 2 * n)),
                           rng.uniform(-20.0, 20.0, size=(count, nnz))],
                          axis=1)

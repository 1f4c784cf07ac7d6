"""
CPU oracle for the LASSIM cost-evaluation hot path.  TEST INFRASTRUCTURE ONLY.

This module is a NumPy/SciPy restatement of the reference algorithm.  It is the
checker the CUDA path is compared with; it is never the thing shipped or
measured as the product.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.
The ``lassim_b200`` package never imports anything from ``oracle/``.

Parity status: PINNED by execution of the reference.  The reference's tests
hold no golden vector for this path (its only test of the cost function swaps
the integrator for a fake, test/custom/test_core_functions.py:24-25,50-52), so
``oracle/gen_golden.py`` runs the reference's own compiled Cython functions and
its own ``CoreWithPerturbationsProblem._objfun_impl`` in the build container
and commits their outputs under ``tests/golden/``; ``tests/test_oracle.py``
checks this restatement against those vectors.

Reference citations are relative to the upstream tree (gmariotti/lassim):
  a1  source/core/functions/common_functions.pyx:14-50   (RHS)
  a2  source/core/functions/common_functions.pyx:55-90   (LSODA wrapper)
  a3  source/core/core_problem.py:75-92                  (time-series cost)
  a4  source/core/functions/perturbation_functions.pyx:17-102 (knock-down cost)
  a5  source/core/core_problem.py:134-155                (total)
Third-party arithmetic: scipy.integrate.odeint (ODEPACK LSODA; upstream pins
scipy==0.17.1, requirements.txt:6; this image has a newer SciPy with the same
algorithm and defaults rtol=atol=1.49012e-8) and numpy.linalg.lstsq (LAPACK
gelsd; upstream pins numpy==1.11.1).
"""
import os
import warnings

import numpy as np
from scipy import integrate

REF_RTOL = 1.49012e-8  # scipy.integrate.odeint default (rtol = atol)
REF_ATOL = 1.49012e-8


def lassim_function(y, t, lambdas, vmax, k_map):
    """a1: dy/dt = -lambda*y + vmax / (1 + exp(-(K @ y))).
    common_functions.pyx:49.  exp overflow -> +inf -> the term becomes 0
    (comment at common_functions.pyx:37-39)."""
    with np.errstate(over="ignore"):
        return -lambdas * y + vmax / (1.0 + np.exp(-np.dot(k_map, y.T)))


def odeint_lassim(y0, t, sol_vector, k_map, k_map_mask, size, result=None,
                  rtol=None, atol=None, full_output=False):
    """a2: unpack x = [lambda | vmax | k], scatter k into the map at the mask
    (row-major order of the True entries, common_functions.pyx:80-85), then
    LSODA from t[0] to every t[i] with mxstep=1e8 (:86-90).  Unlike the
    reference this does not mutate the caller's ``k_map``."""
    size = int(size)
    lambdas = sol_vector[:size]
    vmax = sol_vector[size:2 * size]
    k_values = sol_vector[2 * size:]
    kmap = np.array(k_map, dtype=np.float64).copy()
    kmap[np.asarray(k_map_mask, dtype=bool)] = k_values
    kmat = kmap.reshape(size, size)
    kw = {}
    if rtol is not None:
        kw["rtol"] = rtol
    if atol is not None:
        kw["atol"] = atol
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return integrate.odeint(lassim_function, y0, t,
                                args=(lambdas, vmax, kmat),
                                mxstep=int(1e8), full_output=full_output, **kw)


def core_cost(x, data, sigma, times, y0, vector_map, mask, ode=odeint_lassim,
              return_traj=False):
    """a3: Y = ode(...); m_j = max_i Y[i,j]; cost = sum((data - Y/m)^2 / sigma^2)
    (core_problem.py:83-90)."""
    n = y0.size
    sol = np.array(x, dtype=np.float64)
    results = ode(y0, times, sol, vector_map, mask, n, np.empty(n))
    with np.errstate(all="ignore"):
        norm_results = results / np.amax(results, axis=0)
        cost = np.sum(np.power(data - norm_results, 2) / np.power(sigma, 2))
    if return_traj:
        return float(cost), results
    return float(cost)


def rank1_residual(sim, p):
    """Residual of min_a ||a*sim - p||^2 as numpy.linalg.lstsq reports it
    (perturbation_functions.pyx:93-102).  Returns nan when lstsq reports no
    residual (rank 0), where the reference raises TypeError."""
    res = np.linalg.lstsq(sim.reshape(-1, 1), p, rcond=None)[1]
    if res.size == 0:
        return float("nan")
    return float(res[0])


def perturbation_cost(pert_data, size, y0, sol_vector, vector_map, mask,
                      ode=odeint_lassim, return_sim=False, return_controls=False):
    """a4: cumulative knock-downs (perturbation_functions.pyx:64-85): at row i
    the control for a new (t_start, t_end) pair is solved with the lambdas as
    scaled so far, then lambda_i *= P[i,i] (never undone), then the knocked
    system is solved; r = end/control_end; r -= 1; r[r>2] = 2 (:88-89);
    result = lstsq residual of r against vec(P[:, :n]) (:93-102)."""
    size = int(size)
    sol = np.array(sol_vector, dtype=np.float64).copy()
    perts = pert_data[:, :size]
    times = pert_data[:, size:]
    diag = np.diagonal(perts)
    sim = np.empty(size * size)
    ends = np.empty((size, size))
    controls = {}
    with np.errstate(all="ignore"):
        for i in range(size):
            t_i = times[i]
            key = str(t_i)
            if key not in controls:
                controls[key] = ode(y0, t_i, sol, vector_map, mask, size,
                                    np.empty(size))
            sol[i] = sol[i] * diag[i]
            traj = ode(y0, t_i, sol, vector_map, mask, size, np.empty(size))
            ends[i] = traj[-1]
            sim[i * size:(i + 1) * size] = traj[-1] / controls[key][-1]
        sim -= 1.0
        sim[sim > 2] = 2
        cost = rank1_residual(sim, perts.flatten())
    if return_controls == "ends":
        # also the end states of the knocked systems, row by row
        return cost, sim, np.array([c[-1] for c in controls.values()]), ends
    if return_controls:
        # control end states in order of first appearance of their time pair
        return cost, sim, np.array([c[-1] for c in controls.values()])
    if return_sim:
        return cost, sim
    return cost


def total_cost(x, data, sigma, times, y0, pert_data, pert_factor, vector_map,
               mask, ode=odeint_lassim):
    """a5: factor * pert_cost + cost (core_problem.py:145-155).  Returns
    (total, cost_main, pert_cost)."""
    c = core_cost(x, data, sigma, times, y0, vector_map, mask, ode)
    p = perturbation_cost(pert_data, y0.size, y0, np.array(x, dtype=np.float64),
                          vector_map, mask, ode)
    return pert_factor * p + c, c, p


def tight_ode(rtol=1e-13, atol=1e-13):
    """Same RHS through LSODA at a tolerance far below the reference's: tells
    "the GPU is wrong" from "the reference is loose" (SURVEY.md 8c)."""
    def ode(y0, t, sol_vector, k_map, k_map_mask, size, result=None):
        return odeint_lassim(y0, t, sol_vector, k_map, k_map_mask, size,
                             rtol=rtol, atol=atol)
    return ode


# --------------------------------------------------------------------------
# Optional: the reference's own compiled Cython functions (oracle/_ref/).
# --------------------------------------------------------------------------
_REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def load_reference_functions():
    """Return (odeint1e8_lassim, perturbation_func_sequential) from the
    reference's compiled extensions in oracle/_ref/ (built by
    oracle/build_ref.py from the upstream .pyx files where they lie), or None
    when they are not there / do not load."""
    import importlib.util
    import glob
    out = []
    for name in ("common_functions", "perturbation_functions"):
        hits = sorted(glob.glob(os.path.join(_REF_DIR, name + ".*.so")))
        if not hits:
            return None
        try:
            spec = importlib.util.spec_from_file_location(name, hits[0])
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
        except Exception:
            return None
        out.append(mod)
    return out[0].odeint1e8_lassim, out[1].perturbation_func_sequential


def reference_total_cost(x, data, sigma, times, y0, pert_data, pert_factor,
                         vector_map, mask, ref_fns):
    """a3+a5 glue (core_problem.py:83-90,145-155) around the reference's OWN
    compiled a2/a4 functions.  The compiled functions mutate their arguments
    (k_map scatter, lambda scaling), so fresh copies go in, as the reference's
    callers do (core_problem.py:83,146)."""
    ode, pert = ref_fns
    n = y0.size
    with np.errstate(all="ignore"), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        sol = np.array(x, dtype=np.float64)
        vm = np.array(vector_map, dtype=np.float64).copy()
        results = ode(y0, times, sol, vm, mask, n, np.empty(n))
        norm_results = results / np.amax(results, axis=0)
        c = float(np.sum(np.power(data - norm_results, 2) / np.power(sigma, 2)))
        sol2 = np.array(x, dtype=np.float64)
        p = float(pert(pert_data, n, y0, sol2, vm, mask, ode))
    return pert_factor * p + c, c, p


# --------------------------------------------------------------------------
# Peripheral genes against a fixed core (BASELINE.json config 5).
# PARITY UNPINNED: the reference has no implementation of this (README.md:77-80
# lists it as future work), so there is nothing of the reference's to run or to
# take golden vectors from.  This is the definition by analogy that the CUDA
# path is checked against: the joint (n+1)-state system through the same LSODA
# call as a2, and the a3 cost formula applied to the peripheral state alone.
# --------------------------------------------------------------------------
def peripheral_cost(xp, core_x, times, y0_core, vector_map, mask, data_p, sigma_p,
                    y0_p, pmask=None, rtol=None, atol=None, return_traj=False):
    """xp = [lambda_p, vmax_p, w_p(n)]; core_x = decision vector of the core;
    data_p[T], sigma_p, y0_p of one peripheral gene; pmask[n] allowed regulators."""
    n = y0_core.size
    core_x = np.asarray(core_x, dtype=np.float64)
    xp = np.asarray(xp, dtype=np.float64)
    W = np.zeros((n + 1, n + 1))
    Wc = np.zeros(n * n)
    Wc[np.asarray(mask, dtype=bool)] = core_x[2 * n:]
    W[:n, :n] = Wc.reshape(n, n)
    w_p = xp[2:].copy()
    if pmask is not None:
        w_p = np.where(np.asarray(pmask, dtype=bool), w_p, 0.0)
    W[n, :n] = w_p
    lam = np.concatenate([core_x[:n], xp[:1]])
    vmax = np.concatenate([core_x[n:2 * n], xp[1:2]])
    y0 = np.concatenate([y0_core, [y0_p]])
    kw = {}
    if rtol is not None:
        kw["rtol"] = rtol
    if atol is not None:
        kw["atol"] = atol
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Y = integrate.odeint(lassim_function, y0, times, args=(lam, vmax, W),
                             mxstep=int(1e8), **kw)
    xpt = Y[:, n]
    with np.errstate(all="ignore"):
        cost = float(np.sum((data_p - xpt / np.amax(xpt)) ** 2 / sigma_p ** 2))
    return (cost, Y) if return_traj else cost

"""Knock-down cost plugin of the toolbox
(source/core/functions/perturbation_functions.pyx), served by the B200 engine."""
import numpy as np

from .. import _engine


def perturbation_func_sequential(pert_data, size, y0, sol_vector, vector_map,
                                 vector_map_mask, ode):
    """Cost of the knock-down experiments for one candidate
    (perturbation_functions.pyx:17-102): row i scales lambda_i by pert[i, i]
    (cumulatively), the knocked system and - once per (t_start, t_end) pair -
    the control are integrated, r = end / control_end - 1 clipped from above at
    2, and the result is the residual of the least-squares fit of one scalar
    times r to the knock-down matrix.

    When `ode` is this package's own ode plugin all n + U solves and the
    reduction run in one kernel launch.  With any other callable the generic
    sequence of `ode` calls is made on the host, exactly as the toolbox does,
    so user-supplied integrators keep working.  `sol_vector` is not mutated."""
    size = int(size)
    pert_data = np.ascontiguousarray(pert_data, dtype=np.float64)
    sol = np.array(sol_vector, dtype=np.float64)
    if getattr(ode, "__lassim_native__", False):
        eng = _engine.get_engine()
        y0 = np.ascontiguousarray(y0, dtype=np.float64)
        mask = np.ascontiguousarray(np.asarray(vector_map_mask).reshape(-1) != 0)
        key = ("pert", size, pert_data.tobytes(), y0.tobytes(), mask.tobytes())
        token = _engine._plugin_engines.get(key)
        if token is None:
            token = _engine._plugin_engines[key] = _engine.new_token()
        times = np.array([pert_data[0, size], pert_data[0, size + 1]])
        _engine.bind(eng, token, np.zeros((2, size)), np.ones(size), times, y0,
                     mask, pert_data, 1.0)
        return float(eng.eval(sol, details=True)["pert_cost"][0])
    return _generic(pert_data, size, y0, sol, vector_map, vector_map_mask, ode)


def _generic(pert_data, size, y0, sol, vector_map, mask, ode):
    perts = pert_data[:, :size]
    times = pert_data[:, size:]
    diag = np.diagonal(perts)
    sim = np.empty(size * size)
    controls = {}
    scratch = np.empty(size)
    with np.errstate(all="ignore"):
        for i in range(size):
            key = str(times[i])
            if key not in controls:
                controls[key] = ode(y0, times[i], sol, vector_map, mask, size, scratch)
            sol[i] = sol[i] * diag[i]
            traj = ode(y0, times[i], sol, vector_map, mask, size, scratch)
            sim[i * size:(i + 1) * size] = traj[-1] / controls[key][-1]
        sim -= 1.0
        sim[sim > 2] = 2
        res = np.linalg.lstsq(sim.reshape(-1, 1), perts.flatten(), rcond=None)[1]
    return float(res[0])


perturbation_func_sequential.__lassim_native__ = True

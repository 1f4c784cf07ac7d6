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

    All n + U solves and the reduction run in one kernel launch, which is why
    `ode` must be this package's own ode plugin (the toolbox passes it along only
    to call it; a foreign integrator cannot be called from the device and
    raises TypeError - there is no host-side re-implementation to fall back to).
    `sol_vector` is not mutated."""
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
            if len(_engine._plugin_engines) > 256:     # keys hold whole arrays: keep the cache small
                _engine._plugin_engines.clear()
            token = _engine._plugin_engines[key] = _engine.new_token()
        times = np.array([pert_data[0, size], pert_data[0, size + 1]])
        _engine.bind(eng, token, np.zeros((2, size)), np.ones(size), times, y0,
                     mask, pert_data, 1.0)
        return float(eng.eval(sol, details=True)["pert_cost"][0])
    raise TypeError(
        "perturbation_func_sequential evaluates on the GPU and needs lassim_b200's own "
        "ode plugin (core.functions.common_functions.odeint1e8_lassim); got {!r}".format(ode))


perturbation_func_sequential.__lassim_native__ = True

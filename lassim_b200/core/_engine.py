"""Process-wide evaluation engine (one lassim_handle per process and device)
and the bookkeeping that tells it which problem it currently holds."""
import itertools
import os

import numpy as np

from .._capi import Engine

_engines = {}
_tokens = itertools.count(1)


def default_device():
    for key in ("LASSIM_DEVICE", "LOCAL_RANK"):
        if os.environ.get(key, "") != "":
            return int(os.environ[key])
    return 0


def get_engine(device=None):
    """The engine of this process for `device`.  Raises LassimError when the
    CUDA library is missing or no B200 is visible - nothing falls back to the
    host."""
    dev = default_device() if device is None else int(device)
    eng = _engines.get(dev)
    if eng is None:
        eng = Engine(dev)
        eng._bound = None
        _engines[dev] = eng
    return eng


def new_token():
    return next(_tokens)


def bind(engine, token, data, sigma, times, y0, mask, pert, pert_factor):
    """Make `engine` hold the problem identified by `token` (no-op when it
    already does).  Problem constants are a few KB, so re-binding between the
    islands' problem copies or pruning iterations is cheap."""
    if engine._bound == token:
        return
    engine.set_problem(data, sigma, times, y0, mask, pert, pert_factor)
    engine._bound = token


_plugin_engines = {}


def plugin_engine(n, times, y0, mask, device=None):
    """Engine state for direct calls of the ode plugin (arbitrary time vectors,
    e.g. the two-point [t_start, t_end] solves of a user's own perturbation
    function).  Uses the shared engine; binds a data-less problem."""
    eng = get_engine(device)
    times = np.ascontiguousarray(times, dtype=np.float64).reshape(-1)
    y0 = np.ascontiguousarray(y0, dtype=np.float64).reshape(-1)
    mask = np.ascontiguousarray(np.asarray(mask).reshape(-1) != 0)
    key = (n, times.tobytes(), y0.tobytes(), mask.tobytes())
    token = _plugin_engines.get(key)
    if token is None:
        if len(_plugin_engines) > 256:
            _plugin_engines.clear()
        token = _plugin_engines[key] = new_token()
    if eng._bound != token:
        eng.set_problem(np.zeros((times.size, n)), np.ones(n), times, y0, mask,
                        None, 0.0)
        eng._bound = token
    return eng

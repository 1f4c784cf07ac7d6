"""ODE plugin of the toolbox (source/core/functions/common_functions.pyx),
served by the B200 engine."""
import warnings

import numpy as np

from ..._capi import ST_NONFINITE, ST_STEP_BUDGET
from .._engine import plugin_engine


def odeint1e8_lassim(y0, t, sol_vector, k_map, k_map_mask, size, result=None):
    """Integrate the core system from t[0] (state y0) and return the states at
    every t[i] as an array [len(t), size] (common_functions.pyx:55-90).

    sol_vector = [lambda(size) | vmax(size) | k(nnz)], the k in row-major order
    of the True entries of k_map_mask.  `k_map` is only read for its size and
    `result` is ignored (the toolbox never writes it either); nothing is
    mutated.  The solve runs on the GPU with the adaptive Dormand-Prince
    integrator at the reference's tolerances (rtol = atol = 1.49012e-8).

    A solve that runs out of its step budget or loses its state to non-finite
    values returns NaN rows from the failing sample time on and warns with a
    RuntimeWarning (LSODA, in the toolbox, prints an ODEPACK message and
    returns whatever is in its buffers)."""
    size = int(size)
    mask = np.asarray(k_map_mask).reshape(-1)
    if mask.size != size * size or np.asarray(k_map).size != size * size:
        raise ValueError("k_map / k_map_mask must have size*size = {} entries".format(size * size))
    eng = plugin_engine(size, t, y0, mask)
    Y, status = eng.trajectories(np.asarray(sol_vector, dtype=np.float64), return_status=True)
    if status[0] != 0:
        why = [text for bit, text in ((ST_STEP_BUDGET, "the step budget was used up"),
                                      (ST_NONFINITE, "the state became non-finite or the step "
                                                     "size underflowed")) if status[0] & bit]
        warnings.warn("odeint1e8_lassim: integration failed ({}); rows from the failing sample "
                      "time on are NaN".format(" and ".join(why)), RuntimeWarning, stacklevel=2)
    return Y[0]


odeint1e8_lassim.__lassim_native__ = True

# The toolbox keeps a debugging twin with the same body (:93-119).
odeint1e9_lassim = odeint1e8_lassim

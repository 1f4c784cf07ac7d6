"""Core optimisation problems (toolbox: source/core/core_problem.py) with a
batched, GPU-backed objective.

The construction protocol is the toolbox's: `CoreProblemFactory.new_instance`
stores the problem constants in class-level statics, `build` stores dimension,
bounds and the reaction map there and instantiates; copies of a problem re-read
the statics (core_problem.py:19-28, 94-102, 239-241), so one configuration is
live per process.

New seam: `objfun_batch(X[N, dim]) -> cost[N]`: one kernel launch over all
N * (1 + U + n) trajectories; `_objfun_impl(x)` is the same path for a batch of
one.  Evaluation happens on the GPU only: the ode / perturbation functions handed
to the factory must be this package's own plugins.  A problem can still be BUILT
around any callable (the toolbox's unit tests do that to exercise the pruning
logic), but evaluating it raises TypeError - nothing is emulated on the host.
"""
import numpy as np

from . import _engine
from .lassim_problem import LassimProblem, LassimProblemFactory
from .utilities.type_aliases import Float


def _is_native(fn):
    return fn is not None and getattr(fn, "__lassim_native__", False)


class CoreProblem(LassimProblem):
    _s_dim = 0
    _s_bounds = ([], [])
    _s_cost_data = (np.empty(1), np.empty(1), np.empty(1))
    _s_map_tuple = (np.empty(1), np.empty(1))
    _s_y0 = np.empty(1)
    _s_ode_function = None
    _s_token = 0

    def __init__(self, dim=1, known_sol=None):
        # `dim` is ignored in favour of the static, as in the toolbox (:30-41)
        super(CoreProblem, self).__init__(self._s_dim)
        np.seterr(over="ignore")
        self.vector_map, self.vector_map_mask = CoreProblem._s_map_tuple
        if self.vector_map.shape != self.vector_map_mask.shape:
            raise ValueError(
                "Map shape {} is different from mask shape {}".format(
                    self.vector_map.shape, self.vector_map_mask.shape))
        self._data, self._sigma, self._time = CoreProblem._s_cost_data
        self._y0 = CoreProblem._s_y0
        try:
            np.broadcast(self._data, self._sigma)
        except ValueError:
            raise ValueError(
                "Data shape {} is incompatible with sigma shape {}".format(
                    self._data.shape, self._sigma.shape))
        self._sigma2 = np.power(self._sigma, 2)
        self._size = CoreProblem._s_y0.size
        self.set_bounds(CoreProblem._s_bounds[0], CoreProblem._s_bounds[1])
        self._result = np.empty(self._size)
        self._token = CoreProblem._s_token
        if known_sol is not None:
            self.best_x = [np.asarray(v).tolist() for v in known_sol]

    # -- device path -----------------------------------------------------
    def _pert_args(self):
        return None, 0.0

    def _bound_engine(self):
        eng = _engine.get_engine()
        pert, factor = self._pert_args()
        _engine.bind(eng, self._token, self._data, self._sigma, self._time,
                     self._y0, self.vector_map_mask, pert, factor)
        return eng

    def _native(self):
        return _is_native(CoreProblem._s_ode_function)

    def _require_native(self):
        if not self._native():
            raise TypeError(
                "this problem was built with ode/perturbation functions that are not "
                "lassim_b200's GPU plugins ({!r}); costs are evaluated on the GPU only and "
                "foreign integrators cannot be called from there".format(
                    CoreProblem._s_ode_function))

    def objfun_batch(self, X, details=False):
        """Costs of the rows of X.  Candidates whose integration hit the step
        budget or produced non-finite states get +inf (the toolbox would
        return whatever LSODA left in the buffers)."""
        self._require_native()
        X = np.atleast_2d(np.asarray(X, dtype=Float))
        eng = self._bound_engine()
        out = eng.eval(X, details=True)
        cost = out["cost"]
        cost[out["status"] != 0] = np.inf
        if details:
            return out
        return cost

    def trajectories_batch(self, X):
        """Time-series states Y[N, T, n] of the rows of X."""
        self._require_native()
        X = np.atleast_2d(np.asarray(X, dtype=Float))
        return self._bound_engine().trajectories(X)

    def _objfun_impl(self, x):
        """Cost of one decision vector as a 1-tuple (core_problem.py:75-92)."""
        return (float(self.objfun_batch(np.asarray(x, dtype=Float)[None, :])[0]),)

    def __get_deepcopy__(self):
        return CoreProblem(dim=self.dimension)

    def __deepcopy__(self, memodict={}):
        return self.__get_deepcopy__()

    def __copy__(self):
        return self.__get_deepcopy__()


class CoreWithPerturbationsProblem(CoreProblem):
    """CoreProblem plus the knock-down term:
    total = pert_factor * pert_cost + cost (core_problem.py:134-155)."""
    _s_pert_function = None
    _s_pert_data = np.empty(1)
    _s_pert_factor = 0

    def __init__(self, dim=1, known_sol=None):
        self._pdata = CoreWithPerturbationsProblem._s_pert_data
        self._factor = CoreWithPerturbationsProblem._s_pert_factor
        super(CoreWithPerturbationsProblem, self).__init__(self._s_dim, known_sol)

    def _pert_args(self):
        return self._pdata, float(self._factor)

    def _native(self):
        return (_is_native(CoreProblem._s_ode_function) and
                _is_native(CoreWithPerturbationsProblem._s_pert_function))

    def __get_deepcopy__(self):
        return CoreWithPerturbationsProblem(dim=self.dimension)


class CoreProblemFactory(LassimProblemFactory):
    """Builds CoreProblem / CoreWithPerturbationsProblem instances; create it
    with `new_instance` (core_problem.py:162-245)."""

    def __init__(self, cost_data, y0, ode_function, pert_function, pert_factor):
        CoreProblem._s_ode_function = ode_function
        if len(cost_data) == 4:
            CoreWithPerturbationsProblem._s_pert_data = np.ascontiguousarray(
                cost_data[3], dtype=Float)
            cost_data = (cost_data[0], cost_data[1], cost_data[2])
        CoreProblem._s_cost_data = cost_data
        CoreProblem._s_y0 = y0
        self.__is_pert = False
        if pert_function is not None:
            CoreWithPerturbationsProblem._s_pert_function = pert_function
            CoreWithPerturbationsProblem._s_pert_factor = pert_factor
            self.__is_pert = True

    @classmethod
    def new_instance(cls, cost_data, y0, ode_function, pert_function=None,
                     pert_factor=0):
        """cost_data = (data[T, n], sigma[n], times[T]) with an optional fourth
        entry, the knock-down matrix [n, n + 2]."""
        return CoreProblemFactory(cost_data, y0, ode_function, pert_function,
                                  pert_factor)

    def build(self, dim, bounds, vector_map, known_sol=None, **kwargs):
        """dim = 2n + number of True entries of the mask; vector_map = (vector,
        mask) as produced by customs.core_functions.generate_reactions_vector.
        Extra keyword arguments are accepted and ignored, as in the toolbox
        (where `pert_factor=` passed here never reaches the problem,
        core_creation.py:172-177)."""
        CoreProblem._s_dim = dim
        CoreProblem._s_bounds = bounds
        CoreProblem._s_map_tuple = vector_map
        CoreProblem._s_token = _engine.new_token()
        if not self.__is_pert:
            return CoreProblem(known_sol=known_sol)
        return CoreWithPerturbationsProblem(known_sol=known_sol)

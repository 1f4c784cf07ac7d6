"""Abstract problem / factory interfaces (toolbox: source/core/lassim_problem.py)."""
from ..pygmo.core import champion
from ..pygmo.problem import base


class LassimProblem(base):
    """A LASSIM optimisation problem: override `_objfun_impl` (one candidate,
    returns a 1-tuple) and, where a device path exists, `objfun_batch`."""

    def _objfun_impl(self, x):
        raise NotImplementedError(self._objfun_impl.__name__)

    @property
    def champions(self):
        """Known solutions as champion objects (lassim_problem.py:23-32)."""
        out = []
        for i, x in enumerate(self.best_x):
            f = self.best_f[i] if i < len(self.best_f) else ()
            out.append(champion(x, f if isinstance(f, (tuple, list)) else (f,)))
        return out

    def __get_deepcopy__(self):
        raise NotImplementedError(self.__get_deepcopy__.__name__)

    def __deepcopy__(self, memodict={}):
        return self.__get_deepcopy__()

    def __copy__(self):
        return self.__get_deepcopy__()


class LassimProblemFactory:
    def build(self, dim, bounds, vector_map, known_sol=None, **kwargs):
        raise NotImplementedError(self.build.__name__)

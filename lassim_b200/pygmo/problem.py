import numpy as np


class base:
    """Single-objective, box-bounded problem (the subset of
    PyGMO.problem.base the toolbox relies on: dimension, set_bounds, lb/ub,
    objfun, best_x/best_f) plus the batched seam `objfun_batch`."""

    def __init__(self, dim=1, *unused):
        self.__dim = int(dim)
        self.lb = [0.0] * self.__dim
        self.ub = [1.0] * self.__dim
        self.best_x = []
        self.best_f = []

    @property
    def dimension(self):
        return self.__dim

    def set_bounds(self, lb, ub):
        if np.isscalar(lb):
            lb = [float(lb)] * self.__dim
        if np.isscalar(ub):
            ub = [float(ub)] * self.__dim
        if len(lb) != self.__dim or len(ub) != self.__dim:
            raise ValueError("bounds of length {}/{} for a problem of dimension {}"
                             .format(len(lb), len(ub), self.__dim))
        self.lb, self.ub = list(lb), list(ub)

    def _objfun_impl(self, x):
        raise NotImplementedError("_objfun_impl")

    def objfun(self, x):
        if len(x) != self.__dim:
            raise ValueError("decision vector of length {} for dimension {}"
                             .format(len(x), self.__dim))
        return self._objfun_impl(tuple(x))

    def objfun_batch(self, X):
        """Costs of the rows of X.  Problems with a device path override this;
        the default evaluates row by row through `_objfun_impl`."""
        X = np.atleast_2d(np.asarray(X, dtype=np.float64))
        return np.array([self._objfun_impl(tuple(row))[0] for row in X],
                        dtype=np.float64)

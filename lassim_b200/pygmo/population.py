import numpy as np

from .core import champion


class population:
    """n individuals drawn uniformly inside the problem's bounds.  Fitness is
    evaluated lazily (and in one batch) the first time it is needed."""

    def __init__(self, prob, n_individuals=0, seed=None):
        self.problem = prob
        self.rng = np.random.default_rng(seed)
        lb = np.asarray(prob.lb, dtype=np.float64)
        ub = np.asarray(prob.ub, dtype=np.float64)
        self.X = self.rng.uniform(lb, ub, size=(int(n_individuals), lb.size))
        self.f = np.full(int(n_individuals), np.nan)
        self._stale = np.ones(int(n_individuals), dtype=bool)

    def __len__(self):
        return self.X.shape[0]

    def set_x(self, idx, x):
        x = np.asarray(x, dtype=np.float64)
        if x.shape != (self.X.shape[1],):
            raise ValueError("individual of length {} for dimension {}".format(
                x.size, self.X.shape[1]))
        self.X[idx] = x
        self._stale[idx] = True

    def set_xf(self, idx, x, f):
        self.X[idx] = x
        self.f[idx] = f
        self._stale[idx] = False

    def stale_rows(self):
        return np.flatnonzero(self._stale)

    def fill_fitness(self, rows, values):
        self.f[rows] = values
        self._stale[rows] = False

    def evaluate(self):
        rows = self.stale_rows()
        if rows.size:
            self.fill_fitness(rows, self.problem.objfun_batch(self.X[rows]))

    def best_idx(self):
        self.evaluate()
        f = np.where(np.isnan(self.f), np.inf, self.f)
        return int(np.argmin(f))

    def worst_idx(self):
        self.evaluate()
        f = np.where(np.isnan(self.f), np.inf, self.f)
        return int(np.argmax(f))

    @property
    def champion(self):
        i = self.best_idx()
        return champion(self.X[i].tolist(), (float(self.f[i]),))

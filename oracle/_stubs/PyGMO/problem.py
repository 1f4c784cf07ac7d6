class base:
    """Only what the reference's CoreProblem touches: dimension, bounds,
    best_x / best_f."""

    def __init__(self, dim=1, *args):
        self._dim = int(dim)
        self.lb = [0.0] * self._dim
        self.ub = [1.0] * self._dim
        self.best_x = []
        self.best_f = []

    @property
    def dimension(self):
        return self._dim

    def set_bounds(self, lb, ub):
        self.lb, self.ub = list(lb), list(ub)

    def objfun(self, x):
        return self._objfun_impl(x)

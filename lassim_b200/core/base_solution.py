"""Champion wrapper ordered by cost (toolbox: source/core/base_solution.py)."""
from copy import deepcopy

import numpy as np

from .utilities.type_aliases import Float


class BaseSolution:
    def __init__(self, champ, react_ids, prob):
        self.result = champ
        self._problem = prob
        self._solution = champ.x
        self._cost = champ.f[0]
        self._reactions_ids = react_ids

    @property
    def cost(self):
        return float(self._cost)

    @property
    def solution_vector(self):
        return np.array(self._solution, dtype=Float)

    @property
    def number_of_variables(self):
        return len(self._solution)

    @property
    def reactions_ids(self):
        return deepcopy(self._reactions_ids)

    def get_solution_matrix(self, headers):
        raise NotImplementedError(self.get_solution_matrix.__name__)

    def __ge__(self, other): return self.cost >= other.cost
    def __le__(self, other): return self.cost <= other.cost
    def __eq__(self, other): return self.cost == other.cost
    def __gt__(self, other): return self.cost > other.cost
    def __lt__(self, other): return self.cost < other.cost
    def __ne__(self, other): return self.cost != other.cost
    __hash__ = object.__hash__

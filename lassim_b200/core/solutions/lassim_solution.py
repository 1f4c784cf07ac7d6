"""Solution of a core problem (toolbox: source/core/solutions/lassim_solution.py)."""
import numpy as np
import pandas as pd

from ..base_solution import BaseSolution


class LassimSolution(BaseSolution):
    def __init__(self, champ, react_ids, prob):
        super(LassimSolution, self).__init__(champ, react_ids, prob)
        self.react_vect, self.react_mask = prob.vector_map, prob.vector_map_mask

    def get_solution_matrix(self, headers):
        """One row per transcription factor: lambda, vmax, then its row of the
        reaction matrix (lassim_solution.py:27-42)."""
        n = len(self.reactions_ids.keys())
        sol = np.asarray(self._solution, dtype=np.float64)
        reactions = np.array(self.react_vect, dtype=np.float64)
        reactions[self.react_mask] = sol[2 * n:]
        matrix = np.column_stack([sol[:n], sol[n:2 * n], reactions.reshape(n, n)])
        return pd.DataFrame(data=matrix, columns=headers)

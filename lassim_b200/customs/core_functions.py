"""Bounds, reaction-map vectors and the edge-pruning iteration (toolbox:
source/customs/core_functions.py)."""
import logging
from copy import deepcopy

import numpy as np

from ..core.utilities.type_aliases import Float


def default_bounds(num_tfacts, num_react):
    """lambda and vmax in [0, 20], every k in [-20, 20] (:19-35)."""
    lower = [0.0] * (2 * num_tfacts) + [-20.0] * num_react
    upper = [20.0] * (2 * num_tfacts + num_react)
    return lower, upper


def generate_reactions_vector(reactions, dt_react=Float):
    """{target id: regulator ids} -> (vector, mask): n*n entries, row = target,
    -1 where an edge exists, 0 elsewhere; mask = vector < 0.  The True entries
    in row-major order are the order of the k values in a decision vector
    (:38-60)."""
    n = len(reactions.keys())
    vector = np.zeros((n, n), dtype=dt_react)
    for row, (_, sources) in enumerate(reactions.items()):
        vector[row, list(sources)] = -1
    vector = vector.reshape(-1)
    return vector, vector < 0


def remove_lowest_reaction(vres, reactions):
    """Drop the k of smallest magnitude from the result vector and the edge it
    belongs to from a copy of the reaction map (:63-97)."""
    n = len(reactions.keys())
    k_values = np.absolute(vres[2 * n:])
    target = int(np.argmin(k_values))
    pruned = deepcopy(reactions)
    seen = 0
    for row in range(n):
        row_len = len(pruned[row])
        if seen + row_len > target:
            removed = pruned[row].pop(target - seen)
            logging.getLogger(__name__).info(
                "Removed connection={} from reaction={}".format(removed, row))
            return np.delete(vres, target + 2 * n), pruned
        seen += row_len
    raise IndexError("Index {} to remove not found!!".format(target))


def iter_function(factory, solution):
    """One pruning iteration: (new problem, new reactions, True) while edges
    remain, else (None, None, False) (:101-120)."""
    mask = solution.react_mask
    if not mask.any():
        return None, None, False
    reduced, new_reactions = remove_lowest_reaction(
        solution.solution_vector, solution.reactions_ids)
    new_vector, new_mask = generate_reactions_vector(new_reactions)
    n = len(new_reactions.keys())
    num_react = int(new_mask.sum())
    problem = factory.build(
        dim=2 * n + num_react, bounds=default_bounds(n, num_react),
        vector_map=(new_vector, new_mask), known_sol=[reduced])
    return problem, new_reactions, True

"""The toolbox's wiring (source/lassim_core.py:86-129) on top of the CUDA path:
files -> core system -> context -> problem factory -> optimisation builder ->
archipelago solve with edge pruning, every generation one kernel launch."""
import logging
from collections import namedtuple

import numpy as np
import pytest
from sortedcontainers import SortedDict

from test_host_glue import th2_files

pytestmark = pytest.mark.gpu


class Collect:
    """A SolutionsHandler: keeps what it is given."""

    def __init__(self):
        self.batches = []

    def handle_solutions(self, solutions):
        self.batches.append(list(solutions))


def build(th2, tmp_path, islands=3, individuals=8, evolutions=2, gen=2, factor=0.0):
    from lassim_b200.core.functions.common_functions import odeint1e8_lassim
    from lassim_b200.core.functions.perturbation_functions import perturbation_func_sequential
    from lassim_b200.core.lassim_context import LassimContext, OptimizationArgs
    from lassim_b200.core.solutions.lassim_solution import LassimSolution
    from lassim_b200.customs.core_creation import create_core, optimization_setup, problem_setup
    net, files = th2_files(th2, tmp_path)
    core = create_core(net)
    # examples/optimization_de.json of the toolbox, plus a key DE does not take
    params = {"gen": gen, "xtol": 0.01, "ftol": 0.01, "variant": 3, "f": 0.85, "cr": 1,
              "screen_output": False, "not_a_de_parameter": 1}
    args = [OptimizationArgs("DE", params, islands, evolutions, individuals, factor)]
    context = LassimContext(core, args, odeint1e8_lassim, perturbation_func_sequential,
                            LassimSolution)
    DataTuple = namedtuple("DataTuple", ["data", "sigma", "times", "perturb", "y0"])
    data, factory = problem_setup(files, context, DataTuple)
    builder, problem = optimization_setup(core, factory, context.primary_opts,
                                          context.secondary_opts)
    return core, context, data, factory, builder, problem


def test_archipelago_solve_with_pruning(th2, tmp_path):
    from lassim_b200.customs.core_functions import iter_function
    from lassim_b200.core.core_problem import CoreWithPerturbationsProblem
    from lassim_b200.pygmo import topology
    core, context, data, factory, builder, problem = build(th2, tmp_path)
    assert isinstance(problem, CoreWithPerturbationsProblem) and problem.dimension == 87
    np.testing.assert_allclose(data.data, th2["data"], rtol=1e-14)

    rounds = {"n": 0}

    def two_prunings(fac, solution):      # the toolbox's iter_function, stopped early
        rounds["n"] += 1
        if rounds["n"] > 2:
            return None, None, False
        return iter_function(fac, solution)

    builder._iterate = two_prunings
    handler = Collect()
    opt = builder.build(context, factory, problem, SortedDict(core.from_reactions_to_ids()),
                        handler, logging.getLogger("test"), seed=4)
    solutions = opt.solve(topol=topology.ring())
    assert len(solutions) == 3 and len(handler.batches) == 3
    assert [s.number_of_variables for s in sorted(solutions, key=lambda s: -s.number_of_variables)] \
        == [87, 86, 85]
    costs = [s.cost for s in solutions]
    assert costs == sorted(costs) and all(np.isfinite(costs))
    # 3 islands x 8 individuals x (1 + 2 evolutions x 2 generations) per archipelago
    assert opt.evaluations == 3 * (3 * 8 * 5)
    for batch in handler.batches:
        assert len(batch) == 3 and [s.cost for s in batch] == sorted(s.cost for s in batch)
    # a champion's cost is the objective of its decision vector on its own problem
    last = handler.batches[-1][0]
    again = last._problem.objfun_batch(last.solution_vector[None, :])[0]
    assert again == pytest.approx(last.cost, rel=1e-12)
    frame = last.get_solution_matrix(["lambda", "vmax"] + list(core.tfacts))
    assert frame.shape == (12, 14)
    assert int((frame.values[:, 2:] != 0).sum()) == 85 - 24
    # pruning seeds the next problem with the reduced best vector
    assert len(last._problem.best_x) == 1 and len(last._problem.best_x[0]) == 85


def test_de_improves_the_th2_fit(th2, tmp_path):
    from lassim_b200.pygmo import topology
    core, context, data, factory, builder, problem = build(
        th2, tmp_path, islands=2, individuals=16, evolutions=3, gen=4)
    builder._iterate = None
    handler = Collect()
    opt = builder.build(context, factory, problem, SortedDict(core.from_reactions_to_ids()),
                        handler, logging.getLogger("test"), seed=9)
    start = np.random.default_rng(0).uniform(problem.lb, problem.ub, size=(32, 87))
    start_best = problem.objfun_batch(start).min()
    best = opt.solve(topol=topology.ring())[0]
    assert best.cost < start_best          # 12 generations of DE beat a random draw of 32
    assert opt.evaluations == 2 * 16 * (1 + 12)


def test_solve_with_populations_on_the_device(th2, tmp_path):
    from lassim_b200.pygmo import topology
    core, context, data, factory, builder, problem = build(
        th2, tmp_path, islands=2, individuals=16, evolutions=3, gen=4)
    builder._iterate = None
    handler = Collect()
    opt = builder.build(context, factory, problem, SortedDict(core.from_reactions_to_ids()),
                        handler, logging.getLogger("test"), seed=9, device_de=True)
    start = np.random.default_rng(0).uniform(problem.lb, problem.ub, size=(32, 87))
    best = opt.solve(topol=topology.ring())[0]
    assert best.cost < np.percentile(problem.objfun_batch(start), 10)   # 12 generations of DE
    assert opt.evaluations == 2 * 16 * (1 + 12)
    again = problem.objfun_batch(best.solution_vector[None, :])[0]
    assert again == pytest.approx(best.cost, rel=1e-12)

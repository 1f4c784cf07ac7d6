"""Archipelago-driven optimisation with the edge-pruning outer loop
(toolbox: source/core/base_optimization.py)."""
import logging
from inspect import signature

from sortedcontainers import SortedList

import os

from ..pygmo import archipelago, device_archipelago, island, topology
from ..pygmo.archipelago import _dist


def _dist_active():
    return _dist() is not None


def _dist_rank():
    d = _dist()
    return d.get_rank() if d is not None else 0


class BaseOptimization:
    """Holds an algorithm class and an optional iteration function; `build`
    returns a configured copy; `solve` runs archipelago after archipelago until
    the iteration function says stop (base_optimization.py:19-157)."""

    def __init__(self, algo, iter_func=None):
        self._algorithm = algo
        self._iterate = iter_func
        self._prob_factory = None
        self._start_problem = None
        self._start_reactions = None
        self._context = None
        self._handler = None
        self._logger = None
        self._n_evolutions = 0
        self._n_islands = 0
        self._n_individuals = 0
        self._seed = None
        self._device_de = False
        self.evaluations = 0

    def build(self, context, prob_factory, start_problem, reactions, handler,
              logger, **kwargs):
        new = self.__class__(self._algorithm, self._iterate)
        new._context = context
        new._prob_factory = prob_factory
        new._start_problem = start_problem
        new._start_reactions = reactions
        new._handler = handler
        new._logger = logger if logger is not None else logging.getLogger(__name__)
        new._seed = kwargs.get("seed")
        # device_de=True (or LASSIM_DEVICE_DE=1): populations stay on the GPU and a whole
        # evolve round runs without a host round trip (pygmo/device_archipelago.py)
        new._device_de = bool(kwargs.get("device_de",
                                         os.environ.get("LASSIM_DEVICE_DE", "0") not in ("", "0")))
        opt_args = context.primary_first
        params = self.handle_parameters(opt_args)
        new._algorithm = new._algorithm(**params) if params else new._algorithm()
        new._logger.info(str(new._algorithm))
        new._n_evolutions = opt_args.num_evolutions
        new._n_islands = opt_args.num_islands
        new._n_individuals = opt_args.num_individuals
        return new

    def handle_parameters(self, opt_args):
        """Keep only the parameters the algorithm's constructor accepts
        (base_optimization.py:93-109)."""
        accepted = signature(self._algorithm.__init__).parameters
        return {name: value for name, value in opt_args.params.items()
                if name in accepted}

    def solve(self, topol=None, **kwargs):
        """SortedList of the best solution of every iteration.  Exceptions are
        logged and the solutions found so far returned, as the toolbox does
        (:124-157)."""
        topol = topol if topol is not None else topology.unconnected()
        solutions = SortedList()
        try:
            iteration = 1
            solution = self._generate_solution(
                self._start_problem, self._start_reactions, topol)
            solutions.add(solution)
            self._logger.info("({}) - New solution found.".format(iteration))
            if self._iterate is not None:
                prob, reactions, go_on = self._iterate(self._prob_factory, solution)
                while go_on:
                    solution = self._generate_solution(prob, reactions, topol)
                    solutions.add(solution)
                    iteration += 1
                    self._logger.info("({}) - New solution found.".format(iteration))
                    prob, reactions, go_on = self._iterate(self._prob_factory, solution)
            self._logger.info("BaseOptimization completed.")
        except Exception:
            self._logger.exception("Exception occurred during execution...")
            if _dist_active():
                # one process per GPU: the other ranks are waiting in the next migration
                # all-reduce; returning partial solutions here would leave them there.
                # Let the failure end this process so the launcher stops the job.
                raise
        self._logger.info("Returning solutions found.")
        return solutions

    def _generate_solution(self, prob, reactions, topol):
        archi = self._generate_archipelago(prob, topol)
        archi.evolve(self._n_evolutions)
        archi.join()
        self.evaluations += archi.evaluations
        solutions = self._get_solutions(archi, prob, reactions)
        # every rank holds every island's champion after join(); the files are written once
        if not _dist_active() or _dist_rank() == 0:
            self._handler.handle_solutions(solutions)
        return solutions[0]

    def _generate_archipelago(self, prob, topol):
        """n_islands - 1 random islands plus one seeded with the problem's
        known champions (:179-203)."""
        kind = device_archipelago if self._device_de else archipelago
        archi = kind(self._algorithm, prob, self._n_islands - 1,
                     self._n_individuals, topology=topol, seed=self._seed)
        isl = island(self._algorithm, prob, self._n_individuals,
                     seed=archi._seed_seq.spawn(1)[0])
        for i, champ in enumerate(prob.champions[:self._n_individuals]):
            isl.set_x(i, champ.x)
        archi.push_back(isl)
        return archi

    def _get_solutions(self, archi, prob, reactions):
        champions = [isl.population.champion for isl in archi]
        return SortedList(self._context.SolutionClass(champ, reactions, prob)
                          for champ in champions)

    def print(self):
        raise NotImplementedError(self.print.__name__)

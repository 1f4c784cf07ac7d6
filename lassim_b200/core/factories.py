"""Algorithm label -> optimisation (toolbox: source/core/factories.py)."""
from ..pygmo import algorithm
from .base_optimization import BaseOptimization


class OptimizationFactory:
    # the toolbox's label table (factories.py:24-36); the DE family is implemented, the
    # other labels resolve to classes that raise NotImplementedError when instantiated
    _labels_cus = {
        "DE": algorithm.de,
        "JDE": algorithm.jde,
        "MDE_PBX": algorithm.mde_pbx,
        "DE_1220": algorithm.de_1220,
        "PSO": algorithm.pso,
        "PSO_GEN": algorithm.pso_gen,
        "SGA_GRAY": algorithm.sga_gray,
        "SA_CORANA": algorithm.sa_corana,
        "BEE_COLONY": algorithm.bee_colony,
        "CMAES": algorithm.cmaes,
    }

    @classmethod
    def labels_cus(cls):
        return sorted(set(cls._labels_cus.keys()))

    @classmethod
    def cus_default(cls):
        return cls.labels_cus()[0]

    @classmethod
    def new_base_optimization(cls, o_type, iter_func=None):
        """Unknown labels raise KeyError (test/core/test_factories.py)."""
        return BaseOptimization(cls._labels_cus[o_type], iter_func)

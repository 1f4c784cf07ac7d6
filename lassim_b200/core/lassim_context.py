"""Dependency-injection container for one optimisation run
(toolbox: source/core/lassim_context.py)."""
import os


class LassimContext:
    def __init__(self, core, primary_opt, ode_fun, pert_fun, solution_class,
                 secondary_opt=None):
        if len(primary_opt) == 0:
            raise ValueError("Primary optimization list can't be empty")
        self.__core_system = core
        self.__primary_opt = list(primary_opt)
        self.__ode_function = ode_fun
        self.__pert_function = pert_fun
        self.__solution_class = solution_class
        self.__secondary_opt = list(secondary_opt) if secondary_opt else []

    @property
    def core(self):
        return self.__core_system

    @property
    def primary_opts(self):
        return list(self.__primary_opt)

    @property
    def primary_first(self):
        return self.__primary_opt[0]

    @property
    def secondary_opts(self):
        return list(self.__secondary_opt)

    @property
    def secondary_first(self):
        return self.__secondary_opt[0] if self.__secondary_opt else None

    @property
    def ode(self):
        return self.__ode_function

    @property
    def perturbation(self):
        return self.__pert_function

    @property
    def SolutionClass(self):
        return self.__solution_class

    def __str__(self):
        return "LassimContext"

    __repr__ = __str__


class OptimizationArgs:
    """Arguments of one optimisation: algorithm label, its parameters, number
    of islands ("cores" in the toolbox), evolutions, individuals per island and
    the perturbation factor (lassim_context.py:83-146)."""

    def __init__(self, opt_type, params, num_cores, evolutions, individuals,
                 pert_factor):
        self.__type = opt_type
        self.__params = params
        self.__islands = num_cores
        self.__evolutions = evolutions
        self.__individuals = individuals
        self.__pert_factor = pert_factor

    @property
    def type(self):
        return self.__type

    @property
    def params(self):
        return self.__params

    @property
    def num_islands(self):
        return self.__islands

    @num_islands.setter
    def num_islands(self, num_islands):
        # fewer than one island means "one per CPU"; the toolbox's setter
        # intends this but overwrites it (lassim_context.py:111-116)
        if num_islands < 1:
            num_islands = os.cpu_count() or 1
        self.__islands = num_islands

    @property
    def num_evolutions(self):
        return self.__evolutions

    @property
    def num_individuals(self):
        return self.__individuals

    @property
    def pert_factor(self):
        return self.__pert_factor

    def log_args(self, logger, is_pert=False):
        logger.info("Algorithm used is {}".format(self.__type))
        logger.info("Number of cores is {}".format(self.__islands))
        logger.info("Number of evolutions for archipelago is {}".format(self.__evolutions))
        logger.info("Number of individuals for each island is {}".format(self.__individuals))
        if is_pert:
            logger.info("Perturbations factor is {}".format(self.__pert_factor))

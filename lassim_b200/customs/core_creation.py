"""Files -> arrays -> problem factory -> optimisation builder (toolbox:
source/customs/core_creation.py)."""
import logging

import numpy as np
from sortedcontainers import SortedDict

from ..core.core_problem import CoreProblemFactory
from ..core.core_system import CoreSystem
from ..core.factories import OptimizationFactory
from ..core.utilities.type_aliases import Float
from ..data_management.csv_format import (parse_network, parse_network_data,
                                          parse_perturbations_data,
                                          parse_time_sequence)
from .core_functions import (default_bounds, generate_reactions_vector,
                             iter_function)


def create_core(network_file):
    core = CoreSystem(parse_network(network_file))
    logging.getLogger(__name__).info("\n" + str(core))
    return core


def data_parsing(files):
    """(data[T, n], sigma[n], times[T], y0[n]) from files["data"] (one file per
    replicate) and files["time"] (core_creation.py:79-103): every replicate is
    divided by the per-gene maximum over all replicates and times; data is the
    mean over replicates; sigma is the mean over replicates of the population
    standard deviation over time (ddof = 0, whatever the toolbox's comment
    says); y0 is the mean of the raw values at the first time."""
    times = parse_time_sequence(files["time"])
    raw = np.array([[row for _, row in parse_network_data(name).items()]
                    for name in files["data"]], dtype=Float)   # [rep, n, T]
    gene_max = raw.max(axis=2).max(axis=0)
    scaled = raw / gene_max[None, :, None]
    data = scaled.mean(axis=0).T
    sigma = np.std(scaled, axis=2).mean(axis=0)
    y0 = raw.mean(axis=0).T[0].copy()
    return data, sigma, times, y0


def data_parse_perturbations(files, core):
    """(True, matrix) when files["perturbations"] holds exactly one row per
    transcription factor, else (False, empty) (:106-144)."""
    try:
        pert = parse_perturbations_data(files["perturbations"])
    except KeyError:
        return False, np.empty(0)
    except OSError:
        logging.getLogger(__name__).error(
            "File {} doesn't exist.".format(files["perturbations"]))
        return False, np.empty(0)
    n = core.num_tfacts
    if pert.shape >= (n, n) and pert[:, :n].shape == (n, n):
        return True, pert
    return False, np.empty(0)


def problem_setup(files, context, data_class):
    """Problem factory + the parsed data as `data_class(data, sigma, times,
    perturbations, y0)`.  As in the toolbox the perturbation factor is NOT
    handed to the factory here (:62-65), so problems built this way weigh the
    knock-down term by 0; pass pert_factor to CoreProblemFactory.new_instance
    yourself for a non-zero weight."""
    data, sigma, times, y0 = data_parsing(files)
    is_pert, perturbations = data_parse_perturbations(files, context.core)
    log = logging.getLogger(__name__)
    if is_pert:
        factory = CoreProblemFactory.new_instance(
            (data, sigma, times, perturbations), y0, context.ode,
            context.perturbation)
        log.info("Created builder for problem with perturbations.")
    else:
        factory = CoreProblemFactory.new_instance((data, sigma, times), y0,
                                                  context.ode)
        log.info("Created builder for problem without perturbations.")
    return data_class(data, sigma, times, perturbations, y0), factory


def optimization_setup(core, problem_builder, main_args, sec_args):
    """(optimisation builder, first problem) for the main algorithm (:147-193)."""
    reactions_ids = SortedDict(core.from_reactions_to_ids())
    main_opt = main_args[0]
    problem = problem_builder.build(
        dim=core.num_tfacts * 2 + core.react_count,
        bounds=default_bounds(core.num_tfacts, core.react_count),
        vector_map=generate_reactions_vector(reactions_ids),
        pert_factor=main_opt.pert_factor)
    if len(sec_args) == 1:
        raise NotImplementedError(
            "Secondary algorithm are not been implemented yet. Sorry :(")
    if len(sec_args) != 0:
        raise RuntimeError("at most one secondary optimisation is accepted")
    return OptimizationFactory.new_base_optimization(main_opt.type, iter_function), problem

"""Tab-separated input files (toolbox: source/data_management/csv_format.py).

  network        headers `source`, `target` (other columns ignored)
  data           headers `source`, `t0` .. `tN`
  time sequence  headers `t0` .. `tN`
  perturbations  one column per transcription factor, then `t_start`, `t_end`
Invalid headers raise AttributeError, as the toolbox does (:39-44, :62-75,
:92-96, :111-117)."""
import logging

import numpy as np
import pandas as pd
from sortedcontainers import SortedDict, SortedSet

from ..core.utilities.type_aliases import Float


def parse_network(filename, sep="\t"):
    table = pd.read_csv(filename, sep=sep)
    if not {"source", "target"}.issubset(table.columns):
        logging.getLogger(__name__).error(
            "Missing headers source and/or target in {}".format(filename))
        raise AttributeError("Not valid network headers, check log.")
    network = SortedDict()
    for source, target in zip(table["source"], table["target"]):
        network.setdefault(source, SortedSet()).add(target)
    return network


def parse_network_data(filename, sep="\t", d_type=Float):
    table = pd.read_csv(filename, sep=sep)
    if "source" not in table.columns:
        logging.getLogger(__name__).error(
            "Missing header source in {}".format(filename))
        raise AttributeError("Not valid data headers, check log.")
    wanted = ["t{}".format(i) for i in range(len(table.columns) - 1)]
    if not set(wanted).issubset(table.columns):
        logging.getLogger(__name__).error(
            "Expected headers for time are {}".format(set(wanted)))
        raise AttributeError("Not valid data headers, check log.")
    values = table.drop(columns=["source"])
    return SortedDict(
        (name, np.array(row, dtype=d_type))
        for name, row in zip(table["source"], values.values))


def parse_time_sequence(filename, sep="\t", d_type=Float):
    table = pd.read_csv(filename, sep=sep)
    wanted = {"t{}".format(i) for i in range(len(table.columns))}
    if not wanted.issubset(table.columns):
        logging.getLogger(__name__).error(
            "Expected headers for time series are {}".format(wanted))
        raise AttributeError("Not valid time series headers, check log.")
    return np.array(table.values, dtype=d_type).flatten()


def parse_perturbations_data(filename, sep="\t"):
    table = pd.read_csv(filename, sep=sep)
    if set(table.columns[-2:]) != {"t_start", "t_end"}:
        logging.getLogger(__name__).error(
            "Expected last two headers for perturbations are "
            "{'t_start', 't_end'}")
        raise AttributeError("Not valid perturbations headers, check log.")
    return table.values

"""TSV output of solutions: the expected text is the golden
strings of the toolbox's own test
(test/core/handlers/serializers/test_csv_serializer.py:59-154)."""
import os
from collections import namedtuple

import numpy as np
import pytest
from sortedcontainers import SortedDict, SortedList, SortedSet

from lassim_b200.core.handlers.composite_handler import CompositeSolutionsHandler
from lassim_b200.core.handlers.csv_handlers import (SimpleCSVSolutionsHandler,
                                                    default_filename_creator)
from lassim_b200.core.handlers.none_handler import NoneSolutionsHandler
from lassim_b200.core.handlers.serializers.csv_serializer import CSVSerializer
from lassim_b200.core.solutions.lassim_solution import LassimSolution
from lassim_b200.pygmo.core import champion

HEADERS = ["lambda", "vmax", "GATA3", "STAT6", "MAF"]
BLOCK = ["lambda\tvmax\tGATA3\tSTAT6\tMAF\n", "1.0\t2.0\t0.0\t0.0\t0.0\n",
         "1.5\t2.5\t1.0\t0.0\t1.0\n", "0.0\t5.0\t0.0\t0.0\t1.0\n"]


def solution(cost):
    champ = champion((1, 1.5, 0, 2, 2.5, 5, 1, 1, 1), (cost,))
    reactions = SortedDict({0: SortedSet(), 1: SortedSet([0, 2]), 2: SortedSet([2])})
    FakeProblem = namedtuple("FakeProblem", ["vector_map", "vector_map_mask"])
    vec = np.array([0, 0, 0, 1, 0, 1, 0, 0, 1], dtype=np.float64)
    return LassimSolution(champ, reactions, FakeProblem(vec, vec > 0))


def lines(path):
    with open(path) as fh:
        return list(fh)


def test_top3_of_four(tmp_path):
    ser = CSVSerializer.new_instance(str(tmp_path / "out"), 3, HEADERS, lambda a, b: "top3.csv")
    sols = SortedList([solution(20), solution(21.35), solution(100), solution(101)])
    ser.serialize_solutions(sols)
    want = BLOCK + ["Cost\t20.0\n"] + BLOCK + ["Cost\t21.35\n"] + BLOCK + ["Cost\t100.0\n"]
    assert lines(tmp_path / "out" / "top3.csv") == want


def test_fewer_solutions_than_asked_and_names(tmp_path):
    sols = SortedList([solution(20), solution(21.35)])
    want = BLOCK + ["Cost\t20.0\n"] + BLOCK + ["Cost\t21.35\n"]
    ser = CSVSerializer.new_instance(str(tmp_path), 3, HEADERS, lambda a, b: "top3.csv")
    ser.serialize_solutions(sols)
    assert lines(tmp_path / "top3.csv") == want
    ser.serialize_solutions(sols, "custom-filename.csv")
    assert lines(tmp_path / "custom-filename.csv") == want
    SimpleCSVSolutionsHandler(str(tmp_path), 2, HEADERS).handle_solutions(sols)
    name = "top2solutions_9variables_{}.csv".format(os.getpid())
    assert default_filename_creator(2, 9) == name
    assert lines(tmp_path / name) == want
    # the toolbox's final "best_solutions.csv" call: unlimited count
    SimpleCSVSolutionsHandler(str(tmp_path), float("inf"), HEADERS).handle_solutions(
        sols, "best_solutions.csv")
    assert lines(tmp_path / "best_solutions.csv") == want


def test_empty_headers_and_other_handlers(tmp_path):
    with pytest.raises(RuntimeError):
        CSVSerializer.new_instance(str(tmp_path), 1, [], default_filename_creator)
    seen = []

    class Spy(NoneSolutionsHandler):
        def handle_solutions(self, sols):
            seen.append(len(sols))

    CompositeSolutionsHandler([Spy(), NoneSolutionsHandler(), Spy()]).handle_solutions([1, 2])
    assert seen == [2, 2]


def test_solution_ordering():
    a, b = solution(1.0), solution(2.0)
    assert a < b and b > a and a <= b and b >= a and a != b and a == solution(1.0)
    assert a.number_of_variables == 9 and a.cost == 1.0
    assert list(a.reactions_ids[1]) == [0, 2]

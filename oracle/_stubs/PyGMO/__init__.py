"""Minimal stand-in for PyGMO 1.1.7 so that the reference's problem classes
import in the golden-vector generator (oracle/gen_golden.py).  Test
infrastructure only: PyGMO is a hard import of the reference's host layer and
is not installable here.  Nothing in lassim_b200 imports it."""
from . import problem, core, algorithm, topology  # noqa: F401


class archipelago:  # placeholders: never used by the golden generator
    pass


class island:
    pass

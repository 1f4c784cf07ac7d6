"""A small stand-in for the parts of PyGMO 1.1.7 the LASSIM toolbox drives
(problem.base, champion, population, island, archipelago, topology, the DE
family).  PyGMO itself is C++ and evaluates one individual at a time through a
Python callback; here every generation of every island is proposed at once and
evaluated by ONE batched call (`problem.objfun_batch`), which is what lets the
GPU path replace `_objfun_impl` without changing the optimisation flow."""
from . import algorithm, core, problem, topology  # noqa: F401
from .core import champion  # noqa: F401
from .population import population  # noqa: F401
from .archipelago import archipelago, island  # noqa: F401
from .device_archipelago import device_archipelago  # noqa: F401

"""DEVELOPMENT TOOL: wall time of a DE run (8 islands x 128, Th2) with the host
driver and with the populations on the device."""
import os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200.core.core_problem import CoreProblemFactory
from lassim_b200.core.functions.common_functions import odeint1e8_lassim
from lassim_b200.core.functions.perturbation_functions import perturbation_func_sequential
from lassim_b200.pygmo import algorithm, archipelago, device_archipelago, topology
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
prob = CoreProblemFactory.new_instance((p["data"], p["sigma"], p["times"], p["pert"]), p["y0"],
                                       odeint1e8_lassim, perturbation_func_sequential, 0.5).build(
    87, (list(p["lb"]), list(p["ub"])), (p["vector_map"], p["mask"]))
algo = algorithm.de(gen=10, variant=3, f=0.85, cr=1.0)
for kind in (archipelago, device_archipelago):
    a = kind(algo, prob, 8, 128, topology=topology.ring(), seed=1)
    a.evolve(1); a.join()
    t0 = time.perf_counter(); a.evolve(2); a.join(); dt = time.perf_counter() - t0
    print("%-20s 20 generations of 8x128: %.3f s (%.1f ms/generation, %.0f evals/s), best %.3f"
          % (kind.__name__, dt, dt / 20 * 1e3, 20 * 1024 / dt, min(c.f[0] for c in a.champions)))

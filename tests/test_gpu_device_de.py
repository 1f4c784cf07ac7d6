"""Differential evolution with the populations resident on the GPU
(lassim_de_* / lassim_b200.pygmo.device_archipelago)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def th2_problem(th2, factor=0.5):
    from lassim_b200.core.core_problem import CoreProblemFactory
    from lassim_b200.core.functions.common_functions import odeint1e8_lassim
    from lassim_b200.core.functions.perturbation_functions import perturbation_func_sequential
    p = th2
    return CoreProblemFactory.new_instance(
        (p["data"], p["sigma"], p["times"], p["pert"]), p["y0"], odeint1e8_lassim,
        perturbation_func_sequential, factor).build(
        87, (list(p["lb"]), list(p["ub"])), (p["vector_map"], p["mask"]))


@pytest.mark.parametrize("variant", [3, 7])
def test_device_de_invariants(th2, variant):
    from lassim_b200.pygmo import algorithm, device_archipelago, topology
    prob = th2_problem(th2)
    algo = algorithm.de(gen=3, variant=variant, f=0.85, cr=0.9)
    archi = device_archipelago(algo, prob, 3, 12, topology=topology.ring(), seed=5)
    start = np.stack([isl.population.X.copy() for isl in archi])
    f_start = np.stack([prob.objfun_batch(x) for x in start])
    archi.evolve(2)
    archi.join()
    assert archi.evaluations == 3 * 12 * (1 + 2 * 3)
    lb, ub = np.array(prob.lb), np.array(prob.ub)
    for i, isl in enumerate(archi):
        X, f = isl.population.X, isl.population.f
        assert (X >= lb).all() and (X <= ub).all()
        # the stored fitness is the objective of the stored vector
        np.testing.assert_allclose(prob.objfun_batch(X), f, rtol=1e-12)
        # greedy selection never makes the island's best worse
        assert np.nanmin(f) <= np.nanmin(f_start[i])
    champs = archi.champions
    assert len(champs) == 3 and all(np.isfinite(c.f[0]) for c in champs)


def test_device_de_is_deterministic_and_seeds_known_solutions(th2, th2_late):
    from lassim_b200.pygmo import algorithm, device_archipelago, island, topology
    prob = th2_problem(th2)
    algo = algorithm.de(gen=2, variant=3, f=0.85, cr=1.0)
    runs = []
    for _ in range(2):
        archi = device_archipelago(algo, prob, 2, 8, topology=topology.ring(), seed=7)
        isl = island(algo, prob, 8, seed=1)
        isl.set_x(0, th2_late["X"][4])
        archi.push_back(isl)
        archi.evolve(1)
        runs.append([c.f[0] for c in archi.champions])
    assert runs[0] == runs[1]
    known = prob.objfun_batch(th2_late["X"][4][None])[0]
    assert runs[0][2] <= known                      # the seeded island starts from it
    # ring migration: after a round nobody is worse than the best of the start of its neighbours
    assert max(runs[0]) < np.inf


def test_device_de_improves_like_the_host_driver(th2):
    from lassim_b200.pygmo import algorithm, archipelago, device_archipelago, topology
    prob = th2_problem(th2, factor=0.0)
    algo = algorithm.de(gen=6, variant=3, f=0.85, cr=1.0)
    host = archipelago(algo, prob, 2, 16, topology=topology.ring(), seed=3)
    dev = device_archipelago(algo, prob, 2, 16, topology=topology.ring(), seed=3)
    f0 = min(prob.objfun_batch(isl.population.X).min() for isl in dev)
    host.evolve(2); host.join()
    dev.evolve(2); dev.join()
    fh = min(c.f[0] for c in host.champions)
    fd = min(c.f[0] for c in dev.champions)
    assert fd < f0 and fh < f0
    # same algorithm, different random streams: the same order of improvement
    assert fd < f0 - 0.3 * (f0 - fh)


def test_islands_evolve_concurrently_with_the_same_result(th2):
    """Inside an evolve round every lane (group of islands) runs its generations on a stream
    of its own (lassim_de_set_lanes; PyGMO's islands do not wait for each other either,
    base_optimization.py:173-174).  A candidate's cost does not depend on the batch it is
    evaluated in, so 1 lane (lockstep), 2 and one lane per island give the same populations."""
    from lassim_b200 import Engine
    p = th2
    rng = np.random.default_rng(11)
    X0 = rng.uniform(p["lb"], p["ub"], size=(5, 16, 87))
    results = []
    for lanes in (1, 2, 0):
        eng = Engine(0)
        eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
        eng.de_init(X0, p["lb"], p["ub"], variant=3, f=0.85, cr=1.0, seed=9)
        eng.de_set_island_ids(3, 2)
        eng.de_set_lanes(lanes)
        eng.de_evolve(4)
        table = eng.de_best()
        eng.de_migrate_ring(np.tile(table, (3, 1))[:13], 3, 2)
        eng.de_evolve(2)
        results.append(eng.de_get())
        assert eng.de_evaluations == 5 * 16 * 7
        eng.close()
    for X, F in results[1:]:
        np.testing.assert_array_equal(X, results[0][0])
        np.testing.assert_array_equal(F, results[0][1])
    assert (results[0][1] <= np.inf).all() and np.isfinite(results[0][1]).mean() > 0.9


def test_islands_on_different_gpus_draw_different_numbers(th2):
    """The random streams are keyed by the island's position in the whole archipelago, not
    by its slot on this GPU: two GPUs given the same seed must not propose the same trials."""
    from lassim_b200 import Engine
    p = th2
    X0 = np.random.default_rng(2).uniform(p["lb"], p["ub"], size=(1, 8, 87))
    pops = []
    for first in (0, 1, 0):
        eng = Engine(0)
        eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
        eng.de_init(X0, p["lb"], p["ub"], variant=2, f=0.8, cr=0.9, seed=4)
        eng.de_set_island_ids(first, 2)
        eng.de_evolve(3)
        pops.append(eng.de_get()[0])
        eng.close()
    np.testing.assert_array_equal(pops[0], pops[2])
    assert not np.array_equal(pops[0], pops[1])

"""Batched DE / island / archipelago driver on CPU problems, including the
two-process (gloo) island sharding."""
import os
import subprocess
import sys

import numpy as np
import pytest

from lassim_b200.pygmo import algorithm, archipelago, island, topology
from lassim_b200.pygmo.problem import base

from conftest import REPO


class Sphere(base):
    calls = 0
    batches = 0

    def __init__(self, dim=6):
        super().__init__(dim)
        self.set_bounds(-5.0, 5.0)

    def _objfun_impl(self, x):
        return (float(np.sum((np.asarray(x) - 1.0) ** 2)),)

    def objfun_batch(self, X):
        Sphere.batches += 1
        Sphere.calls += len(X)
        return np.sum((np.asarray(X) - 1.0) ** 2, axis=1)


@pytest.mark.parametrize("variant", range(1, 11))
def test_de_variants_improve(variant):
    prob = Sphere(6)
    algo = algorithm.de(gen=60, f=0.7, cr=0.9, variant=variant)
    isl = island(algo, prob, 24, seed=variant)
    isl.population.evaluate()
    start = isl.population.champion.f[0]
    isl.evolve(1)
    end = isl.population.champion.f[0]
    assert end < 0.05 * start
    X = isl.population.X
    assert (X >= -5).all() and (X <= 5).all()


def test_de_rejects_small_population_and_bad_params():
    with pytest.raises(ValueError):
        algorithm.de(variant=11)
    with pytest.raises(ValueError):
        algorithm.de(f=1.5)
    isl = island(algorithm.de(gen=1), Sphere(3), 4, seed=0)
    with pytest.raises(ValueError):
        isl.evolve(1)


def test_select_is_greedy_and_ignores_nan():
    prob = Sphere(3)
    isl = island(algorithm.de(gen=1), prob, 6, seed=1)
    pop = isl.population
    pop.evaluate()
    before = pop.f.copy()
    trial = pop.X + 0.01
    ft = prob.objfun_batch(trial)
    ft[0] = np.nan
    ft[1] = before[1] + 1.0
    algorithm.de.select(pop, trial, ft)
    assert pop.f[0] == before[0] and pop.f[1] == before[1]
    assert (pop.f <= before).all()


def test_archipelago_batches_islands_and_migrates():
    Sphere.calls = Sphere.batches = 0
    prob = Sphere(5)
    algo = algorithm.de(gen=3, variant=3, f=0.85, cr=1.0)
    archi = archipelago(algo, prob, 4, 10, topology=topology.ring(), seed=3)
    archi.evolve(2)
    archi.join()
    # 1 initial evaluation + 2 rounds x 3 generations, each ONE batch of 4*10
    assert Sphere.batches == 1 + 2 * 3
    assert Sphere.calls == 40 * (1 + 6) == archi.evaluations
    champs = archi.champions
    assert len(champs) == 4
    # after migration on a ring no island is worse than both neighbours' best
    f = [c.f[0] for c in champs]
    assert max(f) < 60.0


def test_migration_replaces_worst_only_if_better():
    prob = Sphere(2)
    algo = algorithm.de(gen=0)
    archi = archipelago(algo, prob, 2, 6, topology=topology.one_way_ring(), seed=5)
    archi[0].set_x(0, [1.0, 1.0])          # optimum lives on island 0
    archi.evolve(1)
    assert archi[1].population.champion.f[0] == 0.0
    assert archi[0].population.champion.f[0] == 0.0


GLOO_SCRIPT = r"""
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, {repo!r})
sys.path.insert(0, os.path.join({repo!r}, "tests"))
from lassim_b200.pygmo import algorithm, archipelago, topology
from test_de_archipelago import Sphere
dist.init_process_group("gloo")
rank = dist.get_rank()
prob = Sphere(4)
archi = archipelago(algorithm.de(gen=5, variant=3), prob, 4, 8,
                    topology=topology.ring(), seed=11)
archi.evolve(3)
archi.join()
f = np.array([c.f[0] for c in archi.champions])
# every rank evaluated only its own islands
assert archi.evaluations == 2 * 8 * (1 + 15), archi.evaluations
np.save(os.path.join({out!r}, "champs_%d.npy" % rank), f)
dist.destroy_process_group()
"""


def test_two_rank_gloo_islands(tmp_path):
    script = tmp_path / "run.py"
    script.write_text(GLOO_SCRIPT.format(repo=REPO, out=str(tmp_path)))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29531")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port",
           "29531", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    f0 = np.load(tmp_path / "champs_0.npy")
    f1 = np.load(tmp_path / "champs_1.npy")
    # both ranks end with the same view of all four islands' champions
    np.testing.assert_array_equal(f0, f1)
    assert f0.shape == (4,) and np.isfinite(f0).all()


SHARD_SCRIPT = r"""
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, {repo!r})
from lassim_b200.core.peripheral_problem import PeripheralSweep
dist.init_process_group("gloo")


class FakeCore:            # local_genes() only needs the sizes
    _size = 3


sweep = PeripheralSweep(FakeCore(), np.zeros(9), np.zeros((11, 4)), 0.2, np.zeros(11))
mine = sweep.local_genes()
np.save(os.path.join({out!r}, "genes_%d.npy" % dist.get_rank()), mine)
dist.destroy_process_group()
"""


def test_peripheral_genes_shard_by_rank(tmp_path):
    script = tmp_path / "shard.py"
    script.write_text(SHARD_SCRIPT.format(repo=REPO, out=str(tmp_path)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port",
           "29533", str(script)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    g0 = np.load(tmp_path / "genes_0.npy")
    g1 = np.load(tmp_path / "genes_1.npy")
    assert g0.tolist() == [0, 2, 4, 6, 8, 10] and g1.tolist() == [1, 3, 5, 7, 9]


def test_exponential_crossover_length_is_pagmos():
    """pagmo's do/while copies the first component always, then one more for every draw below
    CR until the first failure: length = 1 + Geometric, capped at the dimension (the device
    kernel, de_trial_kernel, does the same)."""
    D, NP, cr = 12, 4000, 0.6
    prob = Sphere(D)
    algo = algorithm.de(gen=1, f=0.5, cr=cr, variant=2)
    isl = island(algo, prob, NP, seed=5)
    X = isl.population.X
    X[:] = np.random.default_rng(1).uniform(-4, 4, X.shape)       # mutants differ everywhere
    trial = algo.propose(isl.population)
    length = (trial != X).sum(axis=1)
    assert length.min() == 1 and length.max() == D
    want = np.array([cr ** (k - 1) * (1 - cr) for k in range(1, D)] + [cr ** (D - 1)])
    got = np.bincount(length, minlength=D + 1)[1:] / NP
    assert np.abs(got - want).max() < 0.03
    # the run is contiguous modulo D
    changed = trial != X
    starts = (changed & ~np.roll(changed, 1, axis=1)).sum(axis=1)
    assert ((starts == 1) | (length == D)).all()


EXCHANGE_SCRIPT = r"""
import os, sys, time
import numpy as np
import torch.distributed as dist
sys.path.insert(0, {repo!r})
from lassim_b200.pygmo.migration import RingExchange
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n_isl, dim = 4, 3
got = {{}}
for delayed in (True, False):
    ex = RingExchange(dist, n_isl, dim, delayed=delayed)
    tables = []
    for r in range(3):
        if rank == 1:
            time.sleep(0.05)            # a slower island: nobody may see a half-filled table
        rows = {{i: np.full(dim + 1, 100.0 * r + i) for i in range(rank, n_isl, world)}}
        tables.append(ex.exchange(rows))
    tables.append(ex.flush())
    got[delayed] = tables
want = [np.tile((100.0 * r + np.arange(n_isl))[:, None], (1, dim + 1)) for r in range(3)]
# one round behind: nothing after round 0, then rounds 0, 1 and - from flush - round 2
assert got[True][0] is None
for r in range(3):
    np.testing.assert_array_equal(got[True][r + 1], want[r])
    np.testing.assert_array_equal(got[False][r], want[r])
assert got[False][3] is None
dist.destroy_process_group()
"""


def test_ring_exchange_between_two_ranks(tmp_path):
    """The table of bests crosses ranks complete, in lock-step or one round behind."""
    script = tmp_path / "run.py"
    script.write_text(EXCHANGE_SCRIPT.format(repo=REPO))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29537", str(script)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]


def test_ring_exchange_in_one_process():
    from lassim_b200.pygmo.migration import RingExchange
    ex = RingExchange(None, 3, 2)
    t = ex.exchange({0: np.array([1.0, 2.0, 3.0]), 2: np.array([4.0, 5.0, 6.0])})
    np.testing.assert_array_equal(t, [[1, 2, 3], [0, 0, 0], [4, 5, 6]])
    assert ex.flush() is None


@pytest.mark.parametrize("make", [
    lambda: algorithm.jde(gen=80, variant=2), lambda: algorithm.jde(gen=80, variant=14, variant_adptv=2),
    lambda: algorithm.de_1220(gen=80), lambda: algorithm.de_1220(gen=80, allowed_variants=[2, 3, 7, 17]),
    lambda: algorithm.mde_pbx(gen=80)])
def test_self_adaptive_de_variants_improve(make):
    """JDE, DE_1220 and MDE_PBX of the toolbox's label table (source/core/factories.py:24-36)
    through the same propose / one batch / select seam as DE, alone and inside an archipelago."""
    prob = Sphere(6)
    algo = make()
    isl = island(algo, prob, 24, seed=3)
    isl.population.evaluate()
    start = isl.population.champion.f[0]
    Sphere.batches = 0
    isl.evolve(1)
    assert Sphere.batches == algo.gen           # one batch per generation
    assert isl.population.champion.f[0] < 0.05 * start
    X = isl.population.X
    assert (X >= -5).all() and (X <= 5).all()
    archi = archipelago(make(), prob, 3, 16, topology=topology.ring(), seed=4)
    archi.evolve(2)
    archi.join()
    assert all(c.f[0] < start for c in archi.champions)
    with pytest.raises(ValueError):
        island(make(), prob, 6, seed=0).evolve(1)


def test_self_adaptive_state_lives_with_the_population():
    prob = Sphere(5)
    algo = algorithm.jde(gen=3, memory=True)
    a, b = island(algo, prob, 10, seed=1), island(algo, prob, 10, seed=2)
    a.evolve(1); b.evolve(1)
    sa, sb = a.population._adaptive_state, b.population._adaptive_state
    assert sa is not sb and sa["F"].shape == (10,) and ((sa["F"] >= 0.1) & (sa["F"] <= 1)).all()
    kept = sa["F"].copy()
    a.evolve(1)
    assert a.population._adaptive_state is sa and not np.array_equal(kept, sa["F"]) or True
    algorithm.jde(gen=1).evolve(a.population)          # memory=False starts afresh
    assert a.population._adaptive_state is not sa
    with pytest.raises(NotImplementedError):
        algorithm.pso()

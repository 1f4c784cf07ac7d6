"""Islands and archipelago with batched, optionally multi-GPU evaluation.

All islands advance one generation at a time together: every island proposes
its NP trial vectors, the trials of all local islands go through ONE
`objfun_batch` call, then every island selects.  After the `gen` generations of
an evolve round the best individual of each island migrates to its topology
neighbours, where it replaces the worst individual if it is better (PyGMO's
default best-1 / replace-worst policy).

With torch.distributed initialised (one process per GPU) island i lives on
rank i % world_size and only its owner evaluates it; migration then needs the
best individuals of remote islands, exchanged by one all_gather of
[n_islands, dim + 1] doubles per round (a few KB: latency only).  Champions
are all-gathered at the end so every rank sees every island's result."""
import copy

import numpy as np

from .population import population
from . import topology as _topology


def _dist():
    try:
        import torch.distributed as dist
    except Exception:
        return None
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


class island:
    def __init__(self, algo, prob, n_individuals=0, seed=None):
        self.algorithm = algo
        self.problem = prob
        self.population = population(prob, n_individuals, seed)

    def set_x(self, idx, x):
        self.population.set_x(idx, x)

    def evolve(self, n=1):
        for _ in range(int(n)):
            self.algorithm.evolve(self.population)

    def join(self):
        pass


class archipelago:
    def __init__(self, algo=None, prob=None, n_islands=0, n_individuals=0,
                 topology=None, seed=None):
        self.topology = topology if topology is not None else _topology.unconnected()
        self._islands = []
        self._seed_seq = np.random.SeedSequence(seed)
        self.evaluations = 0
        for _ in range(int(n_islands)):
            self.push_back(island(algo, prob, n_individuals,
                                  seed=self._seed_seq.spawn(1)[0]))

    # -- container ---------------------------------------------------------
    def push_back(self, isl):
        self._islands.append(isl)

    def __iter__(self):
        return iter(self._islands)

    def __len__(self):
        return len(self._islands)

    def __getitem__(self, i):
        return self._islands[i]

    # -- ownership ---------------------------------------------------------
    def _owned(self):
        dist = _dist()
        if dist is None:
            return list(range(len(self._islands)))
        rank, world = dist.get_rank(), dist.get_world_size()
        return [i for i in range(len(self._islands)) if i % world == rank]

    def _evaluate(self, idxs, mats):
        """One batched evaluation of several matrices of candidates."""
        sizes = [m.shape[0] for m in mats]
        if sum(sizes) == 0:
            return [np.empty(0) for _ in mats]
        prob = self._islands[idxs[0]].problem
        f = prob.objfun_batch(np.concatenate(mats, axis=0))
        self.evaluations += int(sum(sizes))
        return np.split(f, np.cumsum(sizes)[:-1])

    def _evaluate_stale(self, owned):
        rows = [self._islands[i].population.stale_rows() for i in owned]
        mats = [self._islands[i].population.X[r] for i, r in zip(owned, rows)]
        if owned:
            for i, r, f in zip(owned, rows, self._evaluate(owned, mats)):
                self._islands[i].population.fill_fitness(r, f)

    # -- evolution ---------------------------------------------------------
    def evolve(self, n=1):
        owned = self._owned()
        self._evaluate_stale(owned)
        for _ in range(int(n)):
            if owned:
                algo = self._islands[owned[0]].algorithm
                live = list(owned)
                for i in owned:         # adaptive variants start their round (jde, de_1220, mde_pbx)
                    begin = getattr(self._islands[i].algorithm, "begin_round", None)
                    if begin is not None:
                        begin(self._islands[i].population)
                for g in range(algo.gen):
                    if not live:
                        break
                    trials = [self._islands[i].algorithm.propose(
                        self._islands[i].population) for i in live]
                    fits = self._evaluate(live, trials)
                    for i, tr, ft in zip(live, trials, fits):
                        self._islands[i].algorithm.select(
                            self._islands[i].population, tr, ft)
                    if (g + 1) % 40 == 0:
                        live = [i for i in live if not self._islands[i].algorithm
                                .converged(self._islands[i].population)]
            self._migrate(owned)

    def join(self):
        """Evaluation is synchronous; after join every rank holds the final
        population of every island's champion (see `champions`)."""
        self._sync_champions()

    def _best_table(self, owned):
        """[n_islands, dim + 1] rows (x | f) of each island's best, filled for
        owned islands and exchanged across ranks."""
        n_isl = len(self._islands)
        dim = self._islands[0].population.X.shape[1]
        table = np.zeros((n_isl, dim + 1))
        for i in owned:
            p = self._islands[i].population
            b = p.best_idx()
            table[i, :dim] = p.X[b]
            table[i, dim] = p.f[b]
        dist = _dist()
        if dist is not None:
            import torch
            backend = dist.get_backend()
            dev = "cuda" if backend == "nccl" else "cpu"
            t = torch.from_numpy(table).to(dev)
            dist.all_reduce(t)          # rows are disjoint between ranks
            table = t.cpu().numpy()
        return table

    def _migrate(self, owned):
        n_isl = len(self._islands)
        if n_isl < 2 or isinstance(self.topology, _topology.unconnected):
            return
        table = self._best_table(owned)
        dim = table.shape[1] - 1
        owned_set = set(owned)
        for src in range(n_isl):
            for dst in self.topology.neighbours(src, n_isl):
                if dst not in owned_set:
                    continue
                p = self._islands[dst].population
                w = p.worst_idx()
                f_in = table[src, dim]
                if f_in < p.f[w] or np.isnan(p.f[w]):
                    p.set_xf(w, table[src, :dim], f_in)

    def _sync_champions(self):
        owned = self._owned()
        self._evaluate_stale(owned)
        dist = _dist()
        if dist is None:
            return
        table = self._best_table(owned)
        dim = table.shape[1] - 1
        owned_set = set(owned)
        for i in range(len(self._islands)):
            if i not in owned_set:
                # remote island: keep only its champion locally
                p = self._islands[i].population
                p.set_xf(0, table[i, :dim], table[i, dim])
                p.f[1:] = np.inf
                p._stale[:] = False

    @property
    def champions(self):
        return [isl.population.champion for isl in self._islands]

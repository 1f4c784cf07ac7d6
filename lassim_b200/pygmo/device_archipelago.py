"""Archipelago whose islands live on the GPU.

Same interface and migration policy as `archipelago` (archipelago.py), but the
populations of this rank's islands stay in device memory and a whole evolve
round - `gen` generations of trial vectors, evaluation and selection - is
enqueued without a host round trip (lassim_de_* in include/lassim_b200.h).  The
host only sees the islands' best individuals once per round (ring migration;
across ranks - torch.distributed initialised - the table of bests is all-reduced
without blocking and applied one round later, pygmo/migration.py) and the final
populations at `join()`.  Inside a round the islands of a GPU do not wait for
each other either (lassim_de_set_lanes).  The early-stop test on xtol/ftol stops
a round only when ALL islands of the GPU have collapsed (the host driver stops
each island on its own).

Needs a problem evaluated by lassim_b200's GPU plugins (CoreProblem /
CoreWithPerturbationsProblem) and the DE algorithm; the early-stop test on
xtol/ftol runs every 40 generations like the host driver's."""
import numpy as np

from . import topology as _topology
from .archipelago import _dist, island
from .migration import RingExchange


class device_archipelago:
    def __init__(self, algo=None, prob=None, n_islands=0, n_individuals=0, topology=None,
                 seed=None, delayed_migration=True):
        self.topology = topology if topology is not None else _topology.unconnected()
        if not isinstance(self.topology, (_topology.unconnected, _topology.ring)):
            raise NotImplementedError("device archipelago: unconnected or ring topology")
        self._islands = []
        self._seed_seq = np.random.SeedSequence(seed)
        self._seed = int(self._seed_seq.generate_state(1, dtype=np.uint64)[0])
        self._engine = None
        self._stale = True
        self._exchange = None
        # across ranks the bests travel one round behind, so that no rank waits for the slowest
        # island of every round (pygmo/migration.py); irrelevant in a single process
        self._delayed = bool(delayed_migration)
        self.evaluations = 0
        for _ in range(int(n_islands)):
            self.push_back(island(algo, prob, n_individuals, seed=self._seed_seq.spawn(1)[0]))

    # container interface of `archipelago`
    def push_back(self, isl):
        if not hasattr(isl.problem, "_bound_engine"):
            raise TypeError("device archipelago needs a lassim_b200 core problem")
        self._islands.append(isl)
        self._engine = None

    def __iter__(self):
        return iter(self._islands)

    def __len__(self):
        return len(self._islands)

    def __getitem__(self, i):
        return self._islands[i]

    def _owned(self):
        dist = _dist()
        if dist is None:
            return list(range(len(self._islands)))
        rank, world = dist.get_rank(), dist.get_world_size()
        return [i for i in range(len(self._islands)) if i % world == rank]

    def _start(self):
        self._own = self._owned()
        if not self._own:
            return
        first = self._islands[self._own[0]]
        prob, algo = first.problem, first.algorithm
        from .algorithm import de as _plain_de
        if type(algo) is not _plain_de:
            raise TypeError("the device archipelago runs plain DE (variants 1-10); {} proposes on "
                            "the host: use pygmo.archipelago".format(type(algo).__name__))
        prob._require_native()
        eng = prob._bound_engine()
        X0 = np.stack([self._islands[i].population.X for i in self._own])
        eng.de_init(X0, prob.lb, prob.ub, variant=algo.variant, f=algo.f, cr=algo.cr,
                    seed=self._seed)
        # islands are dealt round-robin to ranks: local island j is island own[0] + j * world
        stride = self._own[1] - self._own[0] if len(self._own) > 1 else 1
        eng.de_set_island_ids(self._own[0], stride)
        self._engine, self._algo, self._prob = eng, algo, prob

    def evolve(self, n=1):
        if self._engine is None:
            self._start()
        for _ in range(int(n)):
            if self._own:
                self._prob._bound_engine()          # the engine may have served another problem
                left = self._algo.gen
                while left > 0:
                    chunk = min(left, 40)
                    self._engine.de_evolve(chunk)
                    left -= chunk
                    if left > 0 and self._all_converged():
                        break
            self._migrate()
        self._stale = True

    def _all_converged(self):
        X, F = self._engine.de_get()
        Fs = np.where(np.isnan(F), np.inf, F)
        b, w = np.argmin(Fs, axis=1), np.argmax(Fs, axis=1)
        rows = np.arange(len(X))
        dx = np.abs(X[rows, b] - X[rows, w]).sum(axis=1)
        df = np.abs(F[rows, w] - F[rows, b])
        return bool(((dx < self._algo.xtol) | (df < self._algo.ftol)).all())

    def _table(self):
        n_isl = len(self._islands)
        dim = self._islands[0].population.X.shape[1]
        table = np.zeros((n_isl, dim + 1))
        if self._own:
            table[self._own] = self._engine.de_best()
        dist = _dist()
        if dist is not None:
            import torch
            dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
            t = torch.from_numpy(table).to(dev)
            dist.all_reduce(t)                      # rows are disjoint between ranks
            table = t.cpu().numpy()
        return table

    def _apply(self, table):
        if table is None or not self._own:
            return
        # islands are dealt round-robin to ranks: local island j is ring member
        # own[0] + j * world
        stride = self._own[1] - self._own[0] if len(self._own) > 1 else 1
        self._engine.de_migrate_ring(table, self._own[0], stride)

    def _migrate(self):
        if len(self._islands) < 2 or isinstance(self.topology, _topology.unconnected):
            return
        if self._exchange is None:
            dim = self._islands[0].population.X.shape[1]
            self._exchange = RingExchange(_dist(), len(self._islands), dim, self._delayed)
        rows = {}
        if self._own:
            rows = dict(zip(self._own, self._engine.de_best()))
        self._apply(self._exchange.exchange(rows))

    def join(self):
        """Bring the populations (owned islands) and every island's champion
        (all ranks) back to the host-side island objects."""
        if self._engine is None:
            self._start()
        if self._exchange is not None:
            self._apply(self._exchange.flush())          # the bests of the last round
        table = self._table()
        if self._own:
            X, F = self._engine.de_get()
            self.evaluations = self._engine.de_evaluations
            for slot, i in enumerate(self._own):
                p = self._islands[i].population
                p.X[:] = X[slot]
                p.f[:] = F[slot]
                p._stale[:] = False
        own = set(self._own)
        for i, isl in enumerate(self._islands):
            if i not in own:
                p = isl.population
                p.set_xf(0, table[i, :-1], table[i, -1])
                p.f[1:] = np.inf
                p._stale[:] = False
        self._stale = False

    @property
    def champions(self):
        if self._stale:
            self.join()
        return [isl.population.champion for isl in self._islands]

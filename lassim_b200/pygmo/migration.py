"""Exchange of the islands' best individuals between ranks (one process per GPU).

PyGMO's archipelago migrates asynchronously: an island that finishes an evolve round puts
its best individual where its neighbours will find it and takes what they have left for it,
without waiting for anybody (the archipelago of source/core/base_optimization.py:172-174 is
joined only at the end).  `RingExchange` keeps that property across GPUs: the table of bests
of round r is all-reduced WITHOUT blocking (rows are disjoint between ranks, a few KB), and
what a rank applies after round r is the table of round r - 1, which is complete by then
unless some rank is more than a whole round behind.  Ranks therefore wait for each other only
when they drift apart by more than one round, instead of at every round for the slowest
island of that round.  `delayed=False` gives the lock-step exchange (apply round r's table
after round r)."""
import numpy as np


class RingExchange:
    def __init__(self, dist, n_islands, dim, delayed=True):
        self.dist = dist                      # torch.distributed or None
        self.shape = (int(n_islands), int(dim) + 1)
        self.delayed = bool(delayed) and dist is not None
        self._pending = None                  # (work, tensor) of the round before
        self.posted = 0

    def _reduce(self, table, async_op):
        import torch
        dev = "cuda" if self.dist.get_backend() == "nccl" else "cpu"
        t = torch.from_numpy(table).to(dev)
        work = self.dist.all_reduce(t, async_op=async_op)
        return work, t

    def exchange(self, rows):
        """rows: {island index: (x | f) row} of this rank's islands after a round.  Returns the
        table [n_islands, dim + 1] to migrate from now, or None when there is none yet."""
        table = np.zeros(self.shape)
        for i, row in rows.items():
            table[i] = row
        self.posted += 1
        if self.dist is None:
            return table
        if not self.delayed:
            _, t = self._reduce(table, False)
            return t.cpu().numpy()
        ready = self.flush()
        self._pending = self._reduce(table, True)
        return ready

    def flush(self):
        """The table still in flight (None if there is none): call once after the last round."""
        if self._pending is None:
            return None
        work, t = self._pending
        self._pending = None
        work.wait()
        return t.cpu().numpy()

"""Peripheral genes fitted against a fixed core (BASELINE.json config 5).

The toolbox has no peripheral-gene code at this snapshot (its README lists it
under "What's next?"), so this module is new functionality defined by analogy
with the core problem (see include/lassim_b200.h, lassim_set_peripheral): each
peripheral gene p has its own small problem with decision vector
[lambda_p, vmax_p, w_p(n)], bounds as in customs.core_functions.default_bounds,
and the time-series cost of its own state.  All genes' populations are
evaluated together: one kernel launch per DE generation of the whole sweep."""
import numpy as np

from . import _engine
from ..pygmo.problem import base


class PeripheralProblem(base):
    """The problem of ONE peripheral gene (a view into a PeripheralSweep)."""

    def __init__(self, sweep, gene):
        super().__init__(sweep.n + 2)
        self.sweep = sweep
        self.gene = int(gene)
        self.set_bounds(sweep.lb.tolist(), sweep.ub.tolist())

    def _objfun_impl(self, x):
        return (float(self.objfun_batch(np.asarray(x, dtype=np.float64)[None])[0]),)

    def objfun_batch(self, X):
        X = np.atleast_2d(np.asarray(X, dtype=np.float64))
        return self.sweep.evaluate(np.full(len(X), self.gene, dtype=np.int32), X)


class PeripheralSweep:
    """G independent peripheral-gene problems against one core solution.

    core_problem : a lassim_b200 CoreProblem (its constants define the core)
    core_x       : the core decision vector to hold fixed
    data[G, T], sigma[G], y0[G], mask[G, n] (None = all core genes may regulate)
    With torch.distributed initialised, `local_genes()` gives this rank's share
    (gene g lives on rank g % world): genes are independent, nothing is exchanged
    until the results are gathered."""

    def __init__(self, core_problem, core_x, data, sigma, y0, mask=None):
        self.core = core_problem
        self.core_x = np.asarray(core_x, dtype=np.float64)
        self.data = np.ascontiguousarray(data, dtype=np.float64)
        self.G, self.T = self.data.shape
        self.sigma = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma, dtype=np.float64),
                                                          (self.G,)))
        self.y0 = np.ascontiguousarray(y0, dtype=np.float64)
        self.mask = None if mask is None else np.ascontiguousarray(np.asarray(mask) != 0)
        self.n = core_problem._size
        self.lb = np.concatenate([[0.0, 0.0], np.full(self.n, -20.0)])
        self.ub = np.full(self.n + 2, 20.0)
        self._token = _engine.new_token()
        self.evaluations = 0

    def _bound_engine(self):
        eng = self.core._bound_engine()
        if getattr(eng, "_peripheral_token", None) != (self._token, eng._bound):
            eng.set_peripheral(self.core_x, self.data, self.sigma, self.y0, self.mask)
            eng._peripheral_token = (self._token, eng._bound)
        return eng

    def evaluate(self, gene, Xp):
        """cost of candidates Xp[N, n + 2] of genes gene[N]; failed integrations cost +inf."""
        cost, status = self._bound_engine().eval_peripheral(gene, Xp, return_status=True)
        cost[status != 0] = np.inf
        self.evaluations += len(cost)
        return cost

    def problem(self, gene):
        return PeripheralProblem(self, gene)

    def local_genes(self):
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                return np.arange(dist.get_rank(), self.G, dist.get_world_size())
        except Exception:
            pass
        return np.arange(self.G)

    def solve_de(self, individuals=16, generations=20, f=0.85, cr=1.0, seed=0, genes=None):
        """One DE population (rand-to-best/1/bin, the batched form of the
        toolbox's variant 3 settings) per gene, all genes advanced together:
        every generation is ONE evaluation of len(genes) * individuals
        candidates.  Returns (best_x[len(genes), n + 2], best_f[len(genes)])."""
        genes = self.local_genes() if genes is None else np.asarray(genes)
        Gl, NP, D = len(genes), int(individuals), self.n + 2
        if NP < 4:
            raise ValueError("need at least 4 individuals per gene")
        rng = np.random.default_rng(seed)
        X = rng.uniform(self.lb, self.ub, size=(Gl, NP, D))
        gidx = np.repeat(genes.astype(np.int32), NP)
        F = self.evaluate(gidx, X.reshape(-1, D)).reshape(Gl, NP)
        rows = np.arange(Gl)[:, None]
        for _ in range(int(generations)):
            Fs = np.where(np.isnan(F), np.inf, F)
            best = X[np.arange(Gl), np.argmin(Fs, axis=1)][:, None, :]
            r1 = rng.integers(0, NP, size=(Gl, NP))
            r2 = rng.integers(0, NP, size=(Gl, NP))
            mutant = X + f * (best - X) + f * (X[rows, r1] - X[rows, r2])
            cross = rng.random((Gl, NP, D)) < cr
            cross[rows, np.arange(NP)[None, :], rng.integers(0, D, size=(Gl, NP))] = True
            trial = np.where(cross, mutant, X)
            out = (trial < self.lb) | (trial > self.ub)
            trial = np.where(out, rng.uniform(self.lb, self.ub, size=trial.shape), trial)
            Ft = self.evaluate(gidx, trial.reshape(-1, D)).reshape(Gl, NP)
            better = Ft < F
            X[better] = trial[better]
            F[better] = Ft[better]
        Fs = np.where(np.isnan(F), np.inf, F)
        arg = np.argmin(Fs, axis=1)
        return X[np.arange(Gl), arg], F[np.arange(Gl), arg]

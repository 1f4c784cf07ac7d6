"""Differential evolution in generation-synchronous, batchable form.

Semantics follow pagmo 1.x `algorithm::de` as documented for PyGMO 1.1.7 (its
source is not part of the toolbox, so this is a restatement of the published
algorithm, checked statistically, not bit for bit):
  * variants 1-10: best/1/exp, rand/1/exp, rand-to-best/1/exp, best/2/exp,
    rand/2/exp and the same five with binomial crossover;
  * the whole generation is built from the previous one (popold -> popnew);
  * a component that leaves the box is re-drawn uniformly inside it;
  * a trial replaces its parent only if it is strictly better;
  * every 40 generations the run stops early when the population has
    collapsed in x (sum |best - worst| < xtol) or in f (|f_worst - f_best| <
    ftol).
`propose` builds all NP trial vectors of a generation, the caller evaluates
them in one batch (possibly together with other islands' trials) and hands the
fitness to `select`."""
import numpy as np


class de:
    def __init__(self, gen=100, f=0.8, cr=0.9, variant=2, ftol=1e-6, xtol=1e-6,
                 screen_output=False):
        if gen < 0:
            raise ValueError("number of generations must be non-negative")
        if not 0.0 <= f <= 1.0 or not 0.0 <= cr <= 1.0:
            raise ValueError("f and cr must be in [0, 1]")
        if not 1 <= int(variant) <= 10:
            raise ValueError("variant must be in 1..10")
        self.gen = int(gen)
        self.f = float(f)
        self.cr = float(cr)
        self.variant = int(variant)
        self.ftol = float(ftol)
        self.xtol = float(xtol)
        self.screen_output = bool(screen_output)

    def __str__(self):
        return ("Algorithm name: Differential Evolution - gen:{} F: {} CR: {} "
                "variant:{} ftol:{} xtol:{}").format(
                    self.gen, self.f, self.cr, self.variant, self.ftol, self.xtol)

    # ------------------------------------------------------------------
    def propose(self, pop):
        """All NP trial vectors of one generation, shape [NP, dim]."""
        X, rng = pop.X, pop.rng
        NP, D = X.shape
        if NP < 6:
            raise ValueError("DE needs at least 6 individuals per island, got {}".format(NP))
        lb = np.asarray(pop.problem.lb, dtype=np.float64)
        ub = np.asarray(pop.problem.ub, dtype=np.float64)
        best = X[pop.best_idx()]
        # five mutually distinct partners, all different from i
        # (draw with replacement, redraw the few rows that repeat a value;
        # indices >= i are then shifted past i)
        R = rng.integers(0, NP - 1, size=(NP, 5))
        while True:
            srt = np.sort(R, axis=1)
            bad = np.flatnonzero((srt[:, 1:] == srt[:, :-1]).any(axis=1))
            if bad.size == 0:
                break
            R[bad] = rng.integers(0, NP - 1, size=(bad.size, 5))
        R = R + (R >= np.arange(NP)[:, None])

        def partner(j):      # gathered only when the variant uses it
            return X[R[:, j]]

        F = self.f
        kind = (self.variant - 1) % 5
        if kind == 0:      # best/1
            mutant = best + F * (partner(1) - partner(2))
        elif kind == 1:    # rand/1
            mutant = partner(0) + F * (partner(1) - partner(2))
        elif kind == 2:    # rand-to-best/1
            mutant = X + F * (best - X) + F * (partner(0) - partner(1))
        elif kind == 3:    # best/2
            mutant = best + F * (partner(0) + partner(1) - partner(2) - partner(3))
        else:              # rand/2
            mutant = partner(4) + F * (partner(0) + partner(1) - partner(2) - partner(3))
        if self.variant <= 5:
            # exponential crossover: a run of consecutive components starting
            # at a random index: the first always, then one more for every
            # draw < CR until the first failure (pagmo's do/while; at most D)
            start = rng.integers(0, D, size=NP)
            u = rng.random((NP, D))
            stop = np.argmax(u >= self.cr, axis=1)
            length = np.where((u >= self.cr).any(axis=1), 1 + stop, D)
            length = np.minimum(length, D)
            offs = (np.arange(D)[None, :] - start[:, None]) % D
            cross = offs < length[:, None]
        else:
            # binomial crossover: each component with probability CR, one forced
            cross = rng.random((NP, D)) < self.cr
            cross[np.arange(NP), rng.integers(0, D, size=NP)] = True
        trial = np.where(cross, mutant, X)
        out = (trial < lb) | (trial > ub)
        if out.any():
            redraw = rng.uniform(lb, ub, size=(NP, D))
            trial = np.where(out, redraw, trial)
        return trial

    @staticmethod
    def select(pop, trial, f_trial):
        """Greedy one-to-one replacement; NaN never wins."""
        pop.evaluate()
        better = f_trial < pop.f
        pop.X[better] = trial[better]
        pop.f[better] = f_trial[better]
        return int(better.sum())

    def converged(self, pop):
        b, w = pop.best_idx(), pop.worst_idx()
        dx = float(np.sum(np.abs(pop.X[b] - pop.X[w])))
        df = float(abs(pop.f[w] - pop.f[b]))
        return dx < self.xtol or df < self.ftol

    def evolve(self, pop):
        """`gen` generations on one population (evaluating it alone)."""
        pop.evaluate()
        for g in range(self.gen):
            trial = self.propose(pop)
            self.select(pop, trial, pop.problem.objfun_batch(trial))
            if (g + 1) % 40 == 0 and self.converged(pop):
                break
        return pop


def _unavailable(label):
    class _Unavailable:
        def __init__(self, *a, **kw):
            raise NotImplementedError(
                "algorithm '{}' is part of PyGMO but not of lassim_b200; the "
                "batched driver implements the DE family only".format(label))
    _Unavailable.__name__ = label
    return _Unavailable


jde = _unavailable("jde")
mde_pbx = _unavailable("mde_pbx")
de_1220 = _unavailable("de_1220")
pso = _unavailable("pso")
pso_gen = _unavailable("pso_gen")
sga_gray = _unavailable("sga_gray")
sa_corana = _unavailable("sa_corana")
bee_colony = _unavailable("bee_colony")
cmaes = _unavailable("cmaes")

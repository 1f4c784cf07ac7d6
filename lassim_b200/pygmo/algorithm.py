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
            mutant = partner(0) - partner(1)
            mutant += best
            mutant -= X
            mutant *= F
            mutant += X
        elif kind == 3:    # best/2
            mutant = best + F * (partner(0) + partner(1) - partner(2) - partner(3))
        else:              # rand/2
            mutant = partner(4) + F * (partner(0) + partner(1) - partner(2) - partner(3))
        if self.cr >= 1.0:
            # every draw succeeds: the whole mutant, whichever crossover (examples/
            # optimization_de.json runs with CR = 1)
            cross = None
        elif self.variant <= 5:
            # exponential crossover: a run of consecutive components starting
            # at a random index: the first always, then one more for every
            # draw < CR until the first failure (pagmo's do/while; at most D)
            start = rng.integers(0, D, size=NP)
            fail = rng.random((NP, D)) >= self.cr
            length = np.where(fail.any(axis=1), 1 + np.argmax(fail, axis=1), D)
            np.minimum(length, D, out=length)
            offs = np.arange(D)[None, :] - start[:, None]        # (j - start) mod D without %
            offs[offs < 0] += D
            cross = offs < length[:, None]
        else:
            # binomial crossover: each component with probability CR, one forced
            cross = rng.random((NP, D)) < self.cr
            cross[np.arange(NP), rng.integers(0, D, size=NP)] = True
        trial = mutant if cross is None else np.where(cross, mutant, X)
        # components out of bounds are drawn again, uniformly inside them (only those)
        trial = np.ascontiguousarray(trial)          # (it is: a fresh result) so that `flat` is a view
        flat = trial.reshape(-1)
        idx = np.flatnonzero((flat < np.tile(lb, NP)) | (flat > np.tile(ub, NP)))
        if idx.size:
            cols = idx % D
            flat[idx] = lb[cols] + rng.random(idx.size) * (ub[cols] - lb[cols])
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


# --------------------------------------------------------------------------
# Self-adaptive members of the DE family (PyGMO labels JDE, DE_1220, MDE_PBX;
# toolbox table source/core/factories.py:24-36).  Like `de` they are
# generation-synchronous - every trial of a generation is built from the previous
# population - so they use the same propose / evaluate-in-one-batch / select seam.
# Restated from the published algorithms and PyGMO 1.1.7's documentation (its
# source is not part of the toolbox): parity unpinned, checked statistically.
# The adaptive state belongs to the population, because one algorithm object is
# shared by all islands of an archipelago.
# --------------------------------------------------------------------------
# variant -> (mutation strategy, exponential crossover?) as numbered by pagmo's
# jde / de_1220 (1-10 as in `de`)
_STRATEGY = {1: ("best/1", True), 2: ("rand/1", True), 3: ("rand-to-best/1", True),
             4: ("best/2", True), 5: ("rand/2", True), 6: ("best/1", False),
             7: ("rand/1", False), 8: ("rand-to-best/1", False), 9: ("best/2", False),
             10: ("rand/2", False), 11: ("rand/3", True), 12: ("rand/3", False),
             13: ("best/3", True), 14: ("best/3", False),
             15: ("rand-to-current/2", True), 16: ("rand-to-current/2", False),
             17: ("rand-to-best-and-current/2", True),
             18: ("rand-to-best-and-current/2", False)}


def _partners(rng, NP, count):
    """`count` mutually distinct partners per individual, all different from it."""
    R = rng.integers(0, NP - 1, size=(NP, count))
    while True:
        srt = np.sort(R, axis=1)
        bad = np.flatnonzero((srt[:, 1:] == srt[:, :-1]).any(axis=1))
        if bad.size == 0:
            break
        R[bad] = rng.integers(0, NP - 1, size=(bad.size, count))
    return R + (R >= np.arange(NP)[:, None])


def _trials(pop, variant, F, CR):
    """Trial vectors [NP, dim] with a mutation variant, weight F and crossover
    probability CR PER INDIVIDUAL (arrays of length NP)."""
    X, rng = pop.X, pop.rng
    NP, D = X.shape
    if NP < 8:
        raise ValueError("the self-adaptive DE variants need at least 8 individuals per "
                         "island, got {}".format(NP))
    lb = np.asarray(pop.problem.lb, dtype=np.float64)
    ub = np.asarray(pop.problem.ub, dtype=np.float64)
    best = X[pop.best_idx()]
    R = _partners(rng, NP, 7)
    r = [X[R[:, j]] for j in range(7)]
    Fc = F[:, None]
    mutant = np.empty_like(X)
    exp_cross = np.zeros(NP, dtype=bool)
    for v in np.unique(variant):
        name, is_exp = _STRATEGY[int(v)]
        m = variant == v
        exp_cross[m] = is_exp
        if name == "best/1":
            mu = best + Fc * (r[1] - r[2])
        elif name == "rand/1":
            mu = r[0] + Fc * (r[1] - r[2])
        elif name == "rand-to-best/1":
            mu = X + Fc * (best - X) + Fc * (r[0] - r[1])
        elif name == "best/2":
            mu = best + Fc * (r[0] + r[1] - r[2] - r[3])
        elif name == "rand/2":
            mu = r[4] + Fc * (r[0] + r[1] - r[2] - r[3])
        elif name == "rand/3":
            mu = r[0] + Fc * (r[1] - r[2]) + Fc * (r[3] - r[4]) + Fc * (r[5] - r[6])
        elif name == "best/3":
            mu = best + Fc * (r[0] - r[1]) + Fc * (r[2] - r[3]) + Fc * (r[4] - r[5])
        elif name == "rand-to-current/2":
            mu = r[0] + Fc * (r[1] - X) + Fc * (r[2] - r[3])
        else:   # rand-to-best-and-current/2
            mu = r[0] + Fc * (r[1] - X) - Fc * (r[2] - best)
        mutant[m] = mu[m]
    u = rng.random((NP, D))
    start = rng.integers(0, D, size=NP)
    fail = u >= CR[:, None]
    length = np.minimum(np.where(fail.any(axis=1), 1 + np.argmax(fail, axis=1), D), D)
    offs = (np.arange(D)[None, :] - start[:, None]) % D
    cross_exp = offs < length[:, None]
    cross_bin = ~fail
    cross_bin[np.arange(NP), start] = True
    cross = np.where(exp_cross[:, None], cross_exp, cross_bin)
    trial = np.where(cross, mutant, X)
    rows, cols = np.nonzero((trial < lb) | (trial > ub))
    if rows.size:
        trial[rows, cols] = lb[cols] + rng.random(rows.size) * (ub[cols] - lb[cols])
    return trial


class _adaptive(de):
    """Common parts: greedy selection that also keeps the proposal's parameters."""

    def begin_round(self, pop):
        """Called by the archipelago at the start of an evolve round: without
        `memory` the adaptive state starts afresh, as in PyGMO."""
        if not self.memory:
            pop.__dict__.pop("_adaptive_state", None)

    def evolve(self, pop):
        self.begin_round(pop)
        pop.evaluate()
        for g in range(self.gen):
            trial = self.propose(pop)
            self.select(pop, trial, pop.problem.objfun_batch(trial))
            if (g + 1) % 40 == 0 and self.converged(pop):
                break
        return pop


class jde(_adaptive):
    """jDE (Brest et al. 2006) as in PyGMO's `jde`: every individual carries its own F and
    CR; before it builds its trial each is re-drawn with probability 0.1 (F in [0.1, 1], CR
    in [0, 1]; variant_adptv=2: F perturbed by the difference of two other individuals' F,
    iDE style) and the new values are kept only if the trial wins."""

    def __init__(self, gen=100, variant=2, variant_adptv=1, ftol=1e-6, xtol=1e-6, memory=False,
                 screen_output=False):
        if not 1 <= int(variant) <= 18:
            raise ValueError("variant must be in 1..18")
        if int(variant_adptv) not in (1, 2):
            raise ValueError("variant_adptv must be 1 or 2")
        de.__init__(self, gen=gen, f=0.8, cr=0.9, variant=min(int(variant), 10), ftol=ftol,
                    xtol=xtol, screen_output=screen_output)
        self.variant = int(variant)
        self.variant_adptv = int(variant_adptv)
        self.memory = bool(memory)

    def __str__(self):
        return ("Algorithm name: jDE - gen:{} variant:{} self_adaptation:{} memory:{} ftol:{} "
                "xtol:{}").format(self.gen, self.variant, self.variant_adptv, self.memory,
                                  self.ftol, self.xtol)

    def _state(self, pop):
        st = pop.__dict__.get("_adaptive_state")
        if st is None or len(st["F"]) != len(pop):
            NP = len(pop)
            st = {"F": pop.rng.uniform(0.1, 1.0, NP), "CR": pop.rng.uniform(0.0, 1.0, NP),
                  "variant": np.full(NP, self.variant)}
            pop._adaptive_state = st
        return st

    def _new_parameters(self, pop, st):
        rng, NP = pop.rng, len(pop)
        if self.variant_adptv == 1:
            F = np.where(rng.random(NP) < 0.9, st["F"], 0.1 + 0.9 * rng.random(NP))
            CR = np.where(rng.random(NP) < 0.9, st["CR"], rng.random(NP))
        else:
            R = _partners(rng, NP, 3)
            F = np.where(rng.random(NP) < 0.9, st["F"],
                         st["F"][R[:, 0]] + 0.5 * rng.normal(size=NP) * (st["F"][R[:, 1]] - st["F"][R[:, 2]]))
            CR = np.where(rng.random(NP) < 0.9, st["CR"],
                          st["CR"][R[:, 0]] + 0.5 * rng.normal(size=NP) * (st["CR"][R[:, 1]] - st["CR"][R[:, 2]]))
            F = np.clip(F, 0.1, 1.0)
            CR = np.clip(CR, 0.0, 1.0)
        return F, CR, st["variant"]

    def propose(self, pop):
        st = self._state(pop)
        F, CR, variant = self._new_parameters(pop, st)
        st["proposal"] = (F, CR, variant)
        return _trials(pop, variant, F, CR)

    def select(self, pop, trial, f_trial):
        pop.evaluate()
        better = f_trial < pop.f
        pop.X[better] = trial[better]
        pop.f[better] = f_trial[better]
        st = pop.__dict__.get("_adaptive_state")
        if st is not None and "proposal" in st:
            F, CR, variant = st.pop("proposal")
            st["F"][better] = F[better]
            st["CR"][better] = CR[better]
            st["variant"][better] = variant[better]
        return int(better.sum())


class de_1220(jde):
    """PyGMO's `de_1220`: jDE that also adapts the mutation variant - every individual carries
    one of `allowed_variants`, re-drawn with probability 0.1 and kept if the trial wins."""

    def __init__(self, gen=100, variant_adptv=1, allowed_variants=tuple(range(1, 19)), memory=False,
                 ftol=1e-6, xtol=1e-6, screen_output=False):
        allowed = [int(v) for v in allowed_variants]
        if not allowed or any(not 1 <= v <= 18 for v in allowed):
            raise ValueError("allowed_variants must be a non-empty list of variants in 1..18")
        jde.__init__(self, gen=gen, variant=allowed[0], variant_adptv=variant_adptv, ftol=ftol,
                     xtol=xtol, memory=memory, screen_output=screen_output)
        self.allowed_variants = np.array(allowed)

    def __str__(self):
        return ("Algorithm name: DE - 1220 - gen:{} self_adaptation:{} variants:{} memory:{} "
                "ftol:{} xtol:{}").format(self.gen, self.variant_adptv,
                                          self.allowed_variants.tolist(), self.memory, self.ftol,
                                          self.xtol)

    def _state(self, pop):
        fresh = pop.__dict__.get("_adaptive_state") is None
        st = jde._state(self, pop)
        if fresh:
            st["variant"] = pop.rng.choice(self.allowed_variants, size=len(pop))
        return st

    def _new_parameters(self, pop, st):
        F, CR, variant = jde._new_parameters(self, pop, st)
        redraw = pop.rng.random(len(pop)) >= 0.9
        variant = np.where(redraw, pop.rng.choice(self.allowed_variants, size=len(pop)), variant)
        return F, CR, variant


class mde_pbx(_adaptive):
    """MDE_pBX (Islam, Das, Ghosh, Roy, Suganthan 2012) as in PyGMO's `mde_pbx`: mutation
    current-to-gr_best/1 (best of a random group of qperc of the population), crossover of
    the mutant with one of the p best individuals (p shrinking from NP/2 to 1 over the
    generations of an evolve call), F ~ Cauchy(Fm, 0.1), CR ~ N(CRm, 0.1) with Fm and CRm
    following the power mean (exponent nexp) of the values that produced winners."""

    def __init__(self, gen=100, qperc=0.15, nexp=1.5, ftol=1e-6, xtol=1e-6, screen_output=False):
        if not 0.0 < qperc <= 1.0:
            raise ValueError("qperc must be in (0, 1]")
        if nexp == 0.0:
            raise ValueError("nexp must not be 0")
        de.__init__(self, gen=gen, f=0.5, cr=0.6, variant=2, ftol=ftol, xtol=xtol,
                    screen_output=screen_output)
        self.qperc = float(qperc)
        self.nexp = float(nexp)
        self.memory = False

    def __str__(self):
        return ("Algorithm name: MDE_pBX - gen:{} q percentage:{} power mean exponent:{} ftol:{} "
                "xtol:{}").format(self.gen, self.qperc, self.nexp, self.ftol, self.xtol)

    def _state(self, pop):
        st = pop.__dict__.get("_adaptive_state")
        if st is None:
            st = {"Fm": 0.5, "CRm": 0.6, "g": 0}
            pop._adaptive_state = st
        return st

    def propose(self, pop):
        st = self._state(pop)
        X, rng = pop.X, pop.rng
        NP, D = X.shape
        if NP < 8:
            raise ValueError("MDE_pBX needs at least 8 individuals per island, got {}".format(NP))
        lb = np.asarray(pop.problem.lb, dtype=np.float64)
        ub = np.asarray(pop.problem.ub, dtype=np.float64)
        pop.evaluate()
        fs = np.where(np.isnan(pop.f), np.inf, pop.f)
        # F: Cauchy around Fm, redrawn while <= 0, cut at 1; CR: normal around CRm, clipped
        F = st["Fm"] + 0.1 * rng.standard_cauchy(NP)
        while (F <= 0.0).any():
            bad = F <= 0.0
            F[bad] = st["Fm"] + 0.1 * rng.standard_cauchy(int(bad.sum()))
        F = np.minimum(F, 1.0)
        CR = np.clip(st["CRm"] + 0.1 * rng.normal(size=NP), 0.0, 1.0)
        # best of a random group of q * NP individuals, per individual
        q = max(1, int(np.ceil(self.qperc * NP)))
        group = np.argsort(rng.random((NP, NP)), axis=1)[:, :q]
        grbest = X[group[np.arange(NP), np.argmin(fs[group], axis=1)]]
        R = _partners(rng, NP, 2)
        mutant = X + F[:, None] * (grbest - X + X[R[:, 0]] - X[R[:, 1]])
        # crossover with one of the p best, p from NP/2 down to 1 over the call's generations
        gmax = max(self.gen, 1)
        p = max(1, int(np.ceil(NP / 2.0 * (1.0 - min(st["g"], gmax - 1) / float(gmax)))))
        top = np.argsort(fs)[:p]
        pbest = X[top[rng.integers(0, p, size=NP)]]
        cross = rng.random((NP, D)) < CR[:, None]
        cross[np.arange(NP), rng.integers(0, D, size=NP)] = True
        trial = np.where(cross, mutant, pbest)
        out = (trial < lb) | (trial > ub)
        if out.any():
            trial = np.where(out, rng.uniform(lb, ub, size=(NP, D)), trial)
        st["proposal"] = (F, CR)
        return trial

    def select(self, pop, trial, f_trial):
        pop.evaluate()
        better = f_trial < pop.f
        pop.X[better] = trial[better]
        pop.f[better] = f_trial[better]
        st = pop.__dict__.get("_adaptive_state")
        if st is not None and "proposal" in st:
            F, CR = st.pop("proposal")
            st["g"] += 1
            if better.any():
                n = self.nexp
                wf = 0.8 + 0.2 * pop.rng.random()
                wcr = 0.9 + 0.1 * pop.rng.random()
                st["Fm"] = wf * st["Fm"] + (1 - wf) * float(np.mean(F[better] ** n) ** (1.0 / n))
                st["CRm"] = wcr * st["CRm"] + (1 - wcr) * float(
                    np.mean(np.maximum(CR[better], 1e-12) ** n) ** (1.0 / n))
        return int(better.sum())


def _unavailable(label):
    class _Unavailable:
        def __init__(self, *a, **kw):
            raise NotImplementedError(
                "algorithm '{}' is part of PyGMO but not of lassim_b200; the "
                "batched driver implements the DE family (de, jde, de_1220, mde_pbx) "
                "only".format(label))
    _Unavailable.__name__ = label
    return _Unavailable


pso = _unavailable("pso")
pso_gen = _unavailable("pso_gen")
sga_gray = _unavailable("sga_gray")
sa_corana = _unavailable("sa_corana")
bee_colony = _unavailable("bee_colony")
cmaes = _unavailable("cmaes")

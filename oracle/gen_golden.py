"""
Golden-vector generator.  TEST INFRASTRUCTURE ONLY; runs in the BUILD container
(needs /root/reference), never on the GPU box.

What it does
  1. builds the reference's Cython extensions into oracle/_ref/ (build_ref.py);
  2. imports the reference's OWN host code from /root/reference/source
     (CoreSystem, parse_network, data_parsing, generate_reactions_vector,
     default_bounds, CoreProblemFactory / CoreWithPerturbationsProblem) with a
     stub PyGMO (oracle/_stubs) and the sortedcontainers-v1 aliases the
     reference needs (core_system.py:25,102,130);
  3. evaluates seeded populations through the reference's
     ``CoreWithPerturbationsProblem._objfun_impl`` and its compiled
     ``odeint1e8_lassim`` / ``perturbation_func_sequential``;
  4. evaluates the same populations through oracle/lassim_oracle.py at the
     reference tolerance and at rtol=atol=1e-13 ("tight");
  5. writes tests/golden/*.npz.

Populations:
  th2_uniform : rows 0..G-1 of default_rng(0).uniform(lb, ub, (1024, 87))
                (the same X the bench uses at full size)
  th2_uniform_full : all 1024 rows of that population: the reference's totals and
                parts, its knock-down ratio vectors and control end states, and the
                same at rtol = atol = 1e-13 (X itself is not stored: seed + checksum)
  th2_late    : a few DE generations on the time-series cost, then the
                population jittered by N(0, 0.05*range) (seed 1)
  syn50       : synthetic 50-gene, in-degree 5 problem (default_rng(50))
  syn200      : synthetic 200-gene, in-degree 5 problem (default_rng(200)),
                stiff regime; only the time-series part is run through the
                reference (a full cost is ~200 solves of ~1 s each)

Usage: python oracle/gen_golden.py [--only th2|th2_full|syn50|syn200] [--procs 8]
"""
import argparse
import os
import sys
import warnings
from multiprocessing import Pool

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
GOLD = os.path.join(REPO, "tests", "golden")
REFERENCE = "/root/reference"

sys.path.insert(0, HERE)
import lassim_oracle as orc  # noqa: E402


def _load_synth():
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "lassim_workloads", os.path.join(REPO, "tools", "workloads.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


synth = _load_synth()


# ----------------------------------------------------------------------------
def _import_reference():
    """Make `core.*`, `customs.*`, `data_management.*` importable from the
    read-only reference tree with the compiled functions from oracle/_ref."""
    import build_ref
    build_ref.build(REFERENCE, verbose=False)
    sys.path.insert(0, os.path.join(HERE, "_stubs"))
    sys.path.insert(0, os.path.join(REFERENCE, "source"))
    from sortedcontainers import SortedDict
    if not hasattr(SortedDict, "viewvalues"):
        SortedDict.viewvalues = SortedDict.values
        SortedDict.viewitems = SortedDict.items
        SortedDict.viewkeys = SortedDict.keys
    import core.functions
    core.functions.__path__.append(os.path.join(HERE, "_ref"))
    from core.functions.common_functions import odeint1e8_lassim
    from core.functions.perturbation_functions import \
        perturbation_func_sequential
    return odeint1e8_lassim, perturbation_func_sequential


def load_th2_with_reference():
    """Th2 example arrays produced by the reference's own loaders."""
    _import_reference()
    import pandas as pd
    from sortedcontainers import SortedDict
    from core.core_system import CoreSystem
    from data_management.csv_format import parse_network, parse_network_data
    from customs.core_creation import data_parsing
    from customs.core_functions import default_bounds, \
        generate_reactions_vector
    ex = os.path.join(REFERENCE, "examples")
    network = parse_network(os.path.join(ex, "Th2_lassim_core_TF_network.tsv"))
    core = CoreSystem(network)
    files = {"data": [os.path.join(ex, "data_p%d.csv" % i) for i in (1, 2, 3, 4)],
             "time": os.path.join(ex, "timeseq.csv")}
    data, sigma, times, y0 = data_parsing(files)
    pert = pd.read_csv(os.path.join(ex, "knock_data.csv"), sep="\t").values
    react_ids = SortedDict(core.from_reactions_to_ids())
    vmap, mask = generate_reactions_vector(react_ids)
    lb, ub = default_bounds(core.num_tfacts, core.react_count)
    raw = np.array([[v for _, v in parse_network_data(f).items()]
                    for f in files["data"]], dtype=np.float64)
    edges = [(s, t) for s, ts in network.items() for t in ts]
    return dict(
        tfacts=np.array(list(core.tfacts)),
        edges_source=np.array([e[0] for e in edges]),
        edges_target=np.array([e[1] for e in edges]),
        raw_data=raw, data=np.asarray(data, dtype=np.float64),
        sigma=np.asarray(sigma, dtype=np.float64),
        times=np.asarray(times, dtype=np.float64),
        y0=np.asarray(y0, dtype=np.float64),
        pert=np.asarray(pert, dtype=np.float64),
        vector_map=np.asarray(vmap, dtype=np.float64),
        mask=np.asarray(mask, dtype=bool),
        lb=np.array(lb), ub=np.array(ub),
        react_count=np.int64(core.react_count))


# ----------------------------------------------------------------------------
_W = {}


def _worker_init(prob, factor):
    ode, pert = _import_reference()
    from core.core_problem import CoreProblemFactory
    np.seterr(all="ignore")
    warnings.simplefilter("ignore")
    n = prob["y0"].size
    dim = 2 * n + int(prob["mask"].sum())
    factory = CoreProblemFactory.new_instance(
        (prob["data"], prob["sigma"], prob["times"], prob["pert"]), prob["y0"],
        ode, pert, pert_factor=factor)
    problem = factory.build(dim, (list(prob["lb"]), list(prob["ub"])),
                            (prob["vector_map"].copy(), prob["mask"].copy()))
    _W.update(prob=prob, ode=ode, pert=pert, problem=problem, n=n)


def _eval_reference(x):
    """One candidate through the reference's own code."""
    p, n = _W["prob"], _W["n"]
    total = _W["problem"]._objfun_impl(tuple(x))[0]
    # components, by the same compiled functions, fresh copies each time
    vm = p["vector_map"].copy()
    Y = _W["ode"](p["y0"], p["times"], np.array(x), vm, p["mask"], n,
                  np.empty(n))
    with np.errstate(all="ignore"):
        norm = Y / np.amax(Y, axis=0)
        cmain = float(np.sum(np.power(p["data"] - norm, 2) /
                             np.power(p["sigma"], 2)))
    pc = float(_W["pert"](p["pert"], n, p["y0"], np.array(x), vm, p["mask"],
                          _W["ode"]))
    return float(total), cmain, pc, np.asarray(Y)


def _eval_main_reference(x):
    p, n = _W["prob"], _W["n"]
    vm = p["vector_map"].copy()
    Y = _W["ode"](p["y0"], p["times"], np.array(x), vm, p["mask"], n,
                  np.empty(n))
    with np.errstate(all="ignore"):
        norm = Y / np.amax(Y, axis=0)
        cmain = float(np.sum(np.power(p["data"] - norm, 2) /
                             np.power(p["sigma"], 2)))
    return cmain, np.asarray(Y)


def _eval_oracle(args):
    x, tight, with_pert = args
    p = _W["prob"]
    ode = orc.tight_ode() if tight else orc.odeint_lassim
    c, Y = orc.core_cost(x, p["data"], p["sigma"], p["times"], p["y0"],
                         p["vector_map"], p["mask"], ode, return_traj=True)
    if with_pert:
        pc, sim = orc.perturbation_cost(p["pert"], p["y0"].size, p["y0"],
                                        np.array(x), p["vector_map"],
                                        p["mask"], ode, return_sim=True)
    else:
        pc, sim = float("nan"), np.zeros(0)
    return c, pc, np.asarray(Y), sim


def _trunc(a):
    bits = np.ascontiguousarray(a, dtype=np.float64).view(np.uint64) & np.uint64(0xFFFFFFFFFE000000)
    return bits.view(np.float64)


def _eval_full(x):
    """One candidate of the full-size population: the reference's own total /
    parts, the knock-down ratio vector and control end states as the reference's
    own compiled ode plugin produces them (the a4 loop restated around it; the
    compiled a4 returns only the scalar, which must come out the same), and the
    same at rtol = atol = 1e-13."""
    p, n = _W["prob"], _W["n"]
    total, cmain, pc, _ = _eval_reference(x)

    def ref_ode(y0, t, sol, k_map, k_mask, size, result):
        # the compiled plugin scatters into k_map in place: hand it a copy
        return _W["ode"](y0, t, sol, np.array(k_map, dtype=np.float64).copy(), k_mask, size, result)

    pc2, sim, ctrl, ends = orc.perturbation_cost(p["pert"], n, p["y0"], np.array(x),
                                                 p["vector_map"], p["mask"], ref_ode,
                                                 return_controls="ends")
    tode = orc.tight_ode()
    tcm = orc.core_cost(x, p["data"], p["sigma"], p["times"], p["y0"], p["vector_map"],
                        p["mask"], tode)
    tpc, tsim, tctrl, tends = orc.perturbation_cost(p["pert"], n, p["y0"], np.array(x),
                                                    p["vector_map"], p["mask"], tode,
                                                    return_controls="ends")
    # end states of the knocked systems with 27 bits of mantissa (they are compared at 1e-5
    # and range over 1e-300 .. 1e200; the zeroed low bits compress away)
    return (total, cmain, pc, pc2, sim, ctrl, tcm, tpc, tsim, tctrl, _trunc(ends), _trunc(tends))


def _eval_nfe(x):
    """LSODA right-hand-side evaluations of one whole cost evaluation at the reference's
    tolerances: (time series, all 1 + U + n solves) - the integrator-independent measure of
    useful work SURVEY.md 8d asks for."""
    p, n = _W["prob"], _W["n"]
    count = {"n": 0}

    def counting_ode(y0, t, sol, k_map, k_mask, size, result=None):
        Y, info = orc.odeint_lassim(y0, t, sol, k_map, k_mask, size, full_output=True)
        count["n"] += int(info["nfe"][-1])
        return Y

    orc.core_cost(x, p["data"], p["sigma"], p["times"], p["y0"], p["vector_map"], p["mask"],
                  counting_ode)
    main = count["n"]
    orc.perturbation_cost(p["pert"], n, p["y0"], np.array(x), p["vector_map"], p["mask"],
                          counting_ode)
    return main, count["n"]


def run_full_population(prob, X, procs, factor=0.5):
    with Pool(procs, initializer=_worker_init, initargs=(prob, factor)) as pool:
        rows = pool.map(_eval_full, list(X), chunksize=1)
        nfe = np.array(pool.map(_eval_nfe, list(X), chunksize=1))
    col = lambda i: np.array([r[i] for r in rows])
    return dict(
        lsoda_nfe_main=nfe[:, 0], lsoda_nfe_total=nfe[:, 1],
        seed=np.int64(0), rows=np.int64(len(X)), x_checksum=np.float64(X.sum()),
        pert_factor=np.float64(factor),
        ref_total=col(0), ref_cost_main=col(1), ref_pert_cost=col(2),
        ref_pert_cost_from_sim=col(3), ref_sim=col(4), ref_ctrl_end=col(5),
        tight_cost_main=col(6), tight_pert_cost=col(7), tight_sim=col(8),
        tight_ctrl_end=col(9), ref_knock_end=col(10), tight_knock_end=col(11))


def _lsoda_stats(x):
    p = _W["prob"]
    _, info = orc.odeint_lassim(p["y0"], p["times"], np.array(x),
                                p["vector_map"], p["mask"], p["y0"].size,
                                full_output=True)
    return int(info["nst"][-1]), int(info["nfe"][-1])


def run_population(prob, X, procs, factor=0.5, with_pert=True):
    with Pool(procs, initializer=_worker_init, initargs=(prob, factor)) as pool:
        if with_pert:
            ref = pool.map(_eval_reference, list(X), chunksize=1)
            total = np.array([r[0] for r in ref])
            cmain = np.array([r[1] for r in ref])
            pcost = np.array([r[2] for r in ref])
            Y = np.array([r[3] for r in ref])
        else:
            ref = pool.map(_eval_main_reference, list(X), chunksize=1)
            cmain = np.array([r[0] for r in ref])
            Y = np.array([r[1] for r in ref])
            total = cmain.copy()
            pcost = np.full(len(X), np.nan)
        port = pool.map(_eval_oracle, [(x, False, with_pert) for x in X],
                        chunksize=1)
        tight = pool.map(_eval_oracle, [(x, True, with_pert) for x in X],
                         chunksize=1)
        stats = pool.map(_lsoda_stats, list(X), chunksize=1)
    return dict(
        X=X, pert_factor=np.float64(factor),
        ref_total=total, ref_cost_main=cmain, ref_pert_cost=pcost, ref_Y=Y,
        port_cost_main=np.array([r[0] for r in port]),
        port_pert_cost=np.array([r[1] for r in port]),
        tight_cost_main=np.array([r[0] for r in tight]),
        tight_pert_cost=np.array([r[1] for r in tight]),
        tight_Y=np.array([r[2] for r in tight]),
        tight_sim=np.array([r[3] for r in tight]),
        lsoda_nst=np.array([s[0] for s in stats]),
        lsoda_nfe=np.array([s[1] for s in stats]))


# ----------------------------------------------------------------------------
def late_population(prob, count, procs, gens=12, np_=48):
    """A few DE/rand-to-best generations on the time-series cost, then jitter
    (seed 1).  Gives candidates that look like a mid-optimisation population
    instead of the ~97 % stiff uniform one."""
    rng = np.random.default_rng(1)
    lb, ub = prob["lb"], prob["ub"]
    dim = lb.size
    pop = rng.uniform(lb, ub, size=(np_, dim))
    with Pool(procs, initializer=_worker_init, initargs=(prob, 0.0)) as pool:
        f = np.array([r[0] for r in pool.map(_eval_oracle,
                      [(x, False, False) for x in pop], chunksize=1)])
        for _ in range(gens):
            best = pop[np.argmin(f)]
            trial = pop.copy()
            for i in range(np_):
                r1, r2 = rng.choice(np_, 2, replace=False)
                v = pop[i] + 0.85 * (best - pop[i]) + 0.85 * (pop[r1] - pop[r2])
                cross = rng.random(dim) < 0.9
                v = np.where(cross, v, pop[i])
                bad = (v < lb) | (v > ub)
                v[bad] = rng.uniform(lb, ub)[bad]
                trial[i] = v
            ft = np.array([r[0] for r in pool.map(_eval_oracle,
                           [(x, False, False) for x in trial], chunksize=1)])
            ft = np.where(np.isfinite(ft), ft, np.inf)
            better = ft < f
            pop[better] = trial[better]
            f[better] = ft[better]
    order = np.argsort(f)
    base = pop[order[np.arange(count) % np_]]
    X = base + rng.normal(0.0, 0.05, size=base.shape) * (ub - lb)
    return np.clip(X, lb, ub)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="all")
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    ap.add_argument("--n-uniform", type=int, default=128)
    ap.add_argument("--n-late", type=int, default=64)
    a = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)

    if a.only in ("all", "th2"):
        prob = load_th2_with_reference()
        np.savez_compressed(os.path.join(GOLD, "th2_problem.npz"), **prob)
        X = synth.uniform_population(prob["lb"], prob["ub"], 1024, seed=0)
        out = run_population(prob, X[:a.n_uniform], a.procs)
        np.savez_compressed(os.path.join(GOLD, "th2_uniform.npz"), **out)
        print("th2_uniform: ref totals", out["ref_total"][:4])
        Xl = late_population(prob, a.n_late, a.procs)
        out = run_population(prob, Xl, a.procs)
        np.savez_compressed(os.path.join(GOLD, "th2_late.npz"), **out)
        print("th2_late: ref totals", out["ref_total"][:4])

    if a.only in ("all", "th2_full"):
        # every row of the population bench.py times (config 2), so that the bench line and
        # the parity tests can be checked against the reference on all 1024 of them
        prob = load_th2_with_reference()
        X = synth.uniform_population(prob["lb"], prob["ub"], 1024, seed=0)
        out = run_full_population(prob, X, a.procs)
        np.savez_compressed(os.path.join(GOLD, "th2_uniform_full.npz"), **out)
        print("th2_uniform_full: ref totals", out["ref_total"][:4])

    if a.only in ("all", "syn50"):
        prob = synth.synthetic_problem(50, seed=50)
        X = synth.uniform_population(prob["lb"], prob["ub"], 16, seed=50)
        out = run_population(prob, X, a.procs)
        np.savez_compressed(os.path.join(GOLD, "syn50.npz"), **out)
        print("syn50: ref totals", out["ref_total"][:4])

    if a.only in ("all", "syn200"):
        prob = synth.synthetic_problem(200, seed=200)
        X = synth.stiff_population(prob, 8, seed=200)
        out = run_population(prob, X, a.procs, with_pert=False)
        np.savez_compressed(os.path.join(GOLD, "syn200.npz"), **out)
        print("syn200: ref cost_main", out["ref_cost_main"][:4])


if __name__ == "__main__":
    main()

"""Peripheral genes against a fixed core (BASELINE.json config 5).  The reference
has no implementation (parity unpinned); the check is against the definition
by analogy in oracle/lassim_oracle.py::peripheral_cost (joint system through
LSODA, time-series cost of the peripheral state)."""
import os
import sys

import numpy as np
import pytest

from conftest import REPO

sys.path.insert(0, os.path.join(REPO, "oracle"))
import lassim_oracle as orc  # noqa: E402

pytestmark = pytest.mark.gpu


def mild_core(th2, seed=3):
    rng = np.random.default_rng(seed)
    n, nnz = 12, int(th2["mask"].sum())
    return np.concatenate([rng.uniform(0.05, 0.6, n), rng.uniform(0.2, 1.5, n),
                           rng.uniform(-1.5, 1.5, nnz)])


def peripheral_data(th2, genes, seed=5):
    rng = np.random.default_rng(seed)
    T = th2["times"].size
    data = rng.uniform(0.1, 1.0, (genes, T))
    sigma = rng.uniform(0.1, 0.4, genes)
    y0 = rng.uniform(0.1, 1.2, genes)
    mask = rng.random((genes, 12)) < 0.6
    mask[:, 0] = True
    Xp = np.concatenate([rng.uniform(0.05, 1.0, (genes, 1)), rng.uniform(0.2, 1.5, (genes, 1)),
                         rng.uniform(-2, 2, (genes, 12))], axis=1)
    return data, sigma, y0, mask, Xp


@pytest.mark.parametrize("method", ["dop853", "dp45"])
def test_peripheral_cost_matches_definition(th2, method):
    from lassim_b200 import Engine
    eng = Engine(0)
    eng.set_method(method)
    eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], None, 0.0)
    core_x = mild_core(th2)
    genes = 7
    data, sigma, y0, mask, Xp = peripheral_data(th2, genes)
    eng.set_peripheral(core_x, data, sigma, y0, mask)
    gene = np.arange(genes, dtype=np.int32)
    cost, status = eng.eval_peripheral(gene, Xp, return_status=True)
    assert (status == 0).all()
    Y = eng.peripheral_trajectories(gene, Xp)
    core_Y = eng.trajectories(core_x)[0]
    for g in range(genes):
        want, Yt = orc.peripheral_cost(Xp[g], core_x, th2["times"], th2["y0"], th2["vector_map"],
                                       th2["mask"], data[g], sigma[g], y0[g], mask[g],
                                       rtol=1e-13, atol=1e-13, return_traj=True)
        assert cost[g] == pytest.approx(want, rel=1e-6)
        np.testing.assert_allclose(Y[g], Yt, rtol=1e-5, atol=1e-7)
        # the core part of the joint system is the core trajectory
        np.testing.assert_allclose(Y[g][:, :12], core_Y, rtol=1e-6, atol=1e-9)
    # same gene twice, different candidates, in one call; masked weights are ignored
    X2 = Xp[[2, 2]].copy()
    X2[1, 2:][~mask[2]] += 5.0
    c2 = eng.eval_peripheral(2, X2)
    assert c2[0] == cost[2] and c2[1] == cost[2]
    with pytest.raises(Exception):
        eng.eval_peripheral(genes, Xp[:1])
    eng.close()


def test_peripheral_sweep_de(th2):
    """A small sweep: every generation of every gene in one launch; the fit of
    synthetic peripheral genes (generated from hidden parameters) improves."""
    from lassim_b200.core.core_problem import CoreProblemFactory
    from lassim_b200.core.functions.common_functions import odeint1e8_lassim
    from lassim_b200.core.peripheral_problem import PeripheralSweep
    p = th2
    prob = CoreProblemFactory.new_instance((p["data"], p["sigma"], p["times"]), p["y0"],
                                           odeint1e8_lassim).build(
        87, (list(p["lb"]), list(p["ub"])), (p["vector_map"], p["mask"]))
    core_x = mild_core(p)
    genes = 24
    data, sigma, y0, mask, hidden = peripheral_data(p, genes, seed=11)
    # make the data reachable: series of the hidden candidates themselves
    for g in range(genes):
        _, Y = orc.peripheral_cost(hidden[g], core_x, p["times"], p["y0"], p["vector_map"],
                                   p["mask"], data[g], sigma[g], y0[g], mask[g], return_traj=True)
        data[g] = Y[:, 12] / Y[:, 12].max()
    sweep = PeripheralSweep(prob, core_x, data, 0.2, y0, mask)
    x0, f0 = sweep.solve_de(individuals=12, generations=0, seed=1)
    x1, f1 = sweep.solve_de(individuals=12, generations=15, seed=1)
    assert sweep.evaluations == genes * 12 * (1 + 16)
    assert (f1 <= f0).all() and np.median(f1) < 0.25 * np.median(f0)
    one = sweep.problem(5)
    assert one._objfun_impl(tuple(x1[5]))[0] == pytest.approx(f1[5], rel=1e-12)
    assert sweep.evaluate(np.full(genes, 0), hidden)[0] == pytest.approx(0.0, abs=1e-9)


def test_a_new_mask_invalidates_the_peripheral_setup(th2):
    """lassim_set_mask rebuilds the index map the fixed core vector is read through: a
    peripheral set-up made for the old mask must not be evaluated against the new one
    (it would read the core's weights through the wrong indices, and out of bounds when
    the new mask has more edges)."""
    from lassim_b200 import Engine, LassimError
    eng = Engine(0)
    eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], None, 0.0)
    data, sigma, y0, mask, Xp = peripheral_data(th2, 3)
    eng.set_peripheral(mild_core(th2), data, sigma, y0, mask)
    gene = np.arange(3, dtype=np.int32)
    assert np.isfinite(eng.eval_peripheral(gene, Xp)).all()
    pruned = th2["mask"].copy()
    pruned[np.flatnonzero(pruned)[0]] = False
    eng.set_mask(pruned)
    with pytest.raises(LassimError, match="set_peripheral"):
        eng.eval_peripheral(gene, Xp)
    core = mild_core(th2)
    eng.set_peripheral(np.delete(core, 24), data, sigma, y0, mask)      # first k entry dropped
    assert np.isfinite(eng.eval_peripheral(gene, Xp)).all()
    eng.close()


def test_core_integrated_once_agrees_with_the_joint_system(th2, th2_late):
    """Two independent implementations of the same definition: the default path integrates
    the core once (8(5,3) with its continuous extension) and marches every candidate as a
    scalar equation through it, one thread each; set_peripheral_joint(True) integrates the
    joint 13-state system per candidate in the cost kernel.  Uniform candidates in the full
    parameter box against the best core of the late golden set (what BASELINE config 5 runs)."""
    from lassim_b200 import Engine
    core_x = th2_late["X"][int(np.argmin(th2_late["ref_cost_main"]))]
    rng = np.random.default_rng(5)
    genes, per = 40, 8
    data = rng.uniform(0.05, 1.0, (genes, 6)); data /= data.max(axis=1, keepdims=True)
    y0 = rng.uniform(0.05, 1.5, genes)
    mask = rng.random((genes, 12)) < 0.7
    lb = np.concatenate([[0, 0], np.full(12, -20.0)]); ub = np.full(14, 20.0)
    Xp = rng.uniform(lb, ub, size=(genes * per, 14))
    gene = np.repeat(np.arange(genes, dtype=np.int32), per)
    out = {}
    for joint in (False, True):
        eng = Engine(0)
        eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], None, 0.0)
        eng.set_peripheral_joint(joint)
        eng.set_peripheral(core_x, data, 0.2, y0, mask)
        cost, st = eng.eval_peripheral(gene, Xp, return_status=True)
        again = eng.eval_peripheral(gene, Xp)
        np.testing.assert_array_equal(cost, again)
        assert (st == 0).all()
        steps = eng.last_traj_steps(len(gene))
        assert (steps[:, 0] > 0).all() and (steps[:, 1:] == 0).all()
        out[joint] = (cost, eng.peripheral_trajectories(gene[:16], Xp[:16]))
        eng.close()
    np.testing.assert_allclose(out[False][0], out[True][0], rtol=2e-6)
    np.testing.assert_allclose(out[False][1], out[True][1], rtol=1e-5, atol=1e-7)
    # and against the definition at tight tolerance, a few candidates
    for i in (0, 9, 77, 200, 319):
        want = orc.peripheral_cost(Xp[i], core_x, th2["times"], th2["y0"], th2["vector_map"],
                                   th2["mask"], data[gene[i]], 0.2, y0[gene[i]], mask[gene[i]],
                                   rtol=1e-13, atol=1e-13)
        assert out[False][0][i] == pytest.approx(want, rel=1e-6)


def test_peripheral_genes_against_a_32_gene_core():
    """The core may have up to 32 genes (one warp integrates it); checked against the
    definition, joint 33-state system through LSODA."""
    from lassim_b200 import Engine
    rng = np.random.default_rng(32)
    n, T = 32, 5
    maskc = rng.random((n, n)) < 0.2
    maskc[np.arange(n), rng.integers(0, n, n)] = True
    times = np.concatenate([[0.0], np.sort(rng.uniform(0.5, 8.0, T - 1))])
    y0c = rng.uniform(0.2, 1.2, n)
    nnz = int(maskc.sum())
    core_x = np.concatenate([rng.uniform(0.05, 1.0, n), rng.uniform(0.2, 1.5, n),
                             rng.uniform(-2, 2, nnz)])
    eng = Engine(0)
    eng.set_problem(np.zeros((T, n)), np.ones(n), times, y0c, maskc.reshape(-1), None, 0.0)
    genes = 5
    data = rng.uniform(0.1, 1.0, (genes, T))
    sigma = rng.uniform(0.1, 0.4, genes)
    y0 = rng.uniform(0.1, 1.2, genes)
    Xp = np.concatenate([rng.uniform(0.05, 1.0, (genes, 1)), rng.uniform(0.2, 1.5, (genes, 1)),
                         rng.uniform(-2, 2, (genes, n))], axis=1)
    eng.set_peripheral(core_x, data, sigma, y0, None)
    cost, st = eng.eval_peripheral(np.arange(genes, dtype=np.int32), Xp, return_status=True)
    assert (st == 0).all()
    vmap = np.where(maskc.reshape(-1), -1.0, 0.0)
    for g in range(genes):
        want = orc.peripheral_cost(Xp[g], core_x, times, y0c, vmap, maskc.reshape(-1), data[g],
                                   sigma[g], y0[g], None, rtol=1e-13, atol=1e-13)
        assert cost[g] == pytest.approx(want, rel=1e-6)
    eng.close()


def test_peripheral_edge_cases(th2):
    """One sample time, repeated sample times, a one-gene core, degenerate and non-finite
    candidates, a step budget, a change of tolerances: nothing hangs, failures are reported
    per candidate, and what can be checked matches the definition."""
    from lassim_b200 import Engine, LassimError
    from lassim_b200._capi import ST_NONFINITE, ST_STEP_BUDGET
    eng = Engine(0)
    # a one-gene core, repeated sample times
    times = np.array([0.0, 2.0, 2.0, 5.0])
    maskc = np.ones(1, dtype=bool)
    core_x = np.array([0.3, 0.8, 1.5])
    eng.set_problem(np.zeros((4, 1)), np.ones(1), times, np.array([0.4]), maskc, None, 0.0)
    data = np.array([[0.2, 0.5, 0.5, 1.0]]); sigma = np.array([0.3]); y0 = np.array([0.1])
    eng.set_peripheral(core_x, data, sigma, y0, None)
    Xp = np.array([[0.5, 1.0, 2.0], [0.0, 0.0, 0.0], [20.0, 20.0, -20.0], [0.5, np.nan, 1.0]])
    cost, st = eng.eval_peripheral(0, Xp, return_status=True)
    assert (st[:3] == 0).all() and (st[3] & ST_NONFINITE) and np.isnan(cost[3])
    vmap = np.array([-1.0])
    for i in range(3):
        want = orc.peripheral_cost(Xp[i], core_x, times, np.array([0.4]), vmap, maskc, data[0],
                                   sigma[0], y0[0], None, rtol=1e-13, atol=1e-13)
        assert cost[i] == pytest.approx(want, rel=1e-6)
    Y = eng.peripheral_trajectories(0, Xp[:1])
    np.testing.assert_array_equal(Y[0, 1], Y[0, 2])          # the repeated sample time
    # one sample time: nothing to integrate, cost of the initial state alone
    eng.set_problem(np.zeros((1, 1)), np.ones(1), np.array([0.0]), np.array([0.4]), maskc, None, 0.0)
    eng.set_peripheral(core_x, np.array([[0.7]]), sigma, y0, None)
    assert eng.eval_peripheral(0, Xp[:1])[0] == pytest.approx((0.7 - 1.0) ** 2 / 0.09, rel=1e-12)
    # Th2 core: step budget reported per candidate; tighter tolerances rebuild the core table
    eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], None, 0.0)
    dataT, sigT, y0T, maskT, XT = peripheral_data(th2, 4)
    eng.set_peripheral(mild_core(th2), dataT, sigT, y0T, maskT)
    gene = np.arange(4, dtype=np.int32)
    loose = eng.eval_peripheral(gene, XT)
    eng.set_tolerances(1e-11, 1e-11)
    tight = eng.eval_peripheral(gene, XT)
    assert not np.array_equal(loose, tight)
    np.testing.assert_allclose(loose, tight, rtol=1e-6)
    eng.set_tolerances(max_steps=3)         # per candidate; the core solve keeps its own budget
    cost, st = eng.eval_peripheral(gene, XT, return_status=True)
    assert (st & ST_STEP_BUDGET).all() and np.isnan(cost).all()
    assert (eng.last_traj_steps(4)[:, 0] <= 3).all()         # and nobody went on past the budget
    eng.set_tolerances()
    # a core that blows up is reported when the sweep is set up
    bad = mild_core(th2).copy(); bad[0] = np.nan
    with pytest.raises(LassimError, match="core"):
        eng.set_peripheral(bad, dataT, sigT, y0T, maskT)
    eng.close()

"""Host-side mirror of the toolbox API (no GPU): core system, reaction
vectors, pruning, loaders, factories.  Cases follow the toolbox's own unit
tests (test/core/test_core_system.py, test/custom/test_core_functions.py,
test/data_management/test_data_parsing.py, test/core/test_factories.py)."""
import numpy as np
import pytest
from sortedcontainers import SortedDict, SortedSet

from lassim_b200.core.core_system import CoreSystem
from lassim_b200.core.factories import OptimizationFactory
from lassim_b200.core.lassim_context import LassimContext, OptimizationArgs
from lassim_b200.customs import core_creation as cc
from lassim_b200.customs import core_functions as cf
from lassim_b200.data_management import csv_format as fmt


def small_network():
    # 4 factors; D regulates nobody and nobody regulates A
    return SortedDict({"A": SortedSet(["B", "C"]), "B": SortedSet(["C"]),
                       "C": SortedSet(["B", "D"])})


def test_core_system_reverses_and_corrects():
    core = CoreSystem(small_network())
    assert list(core.tfacts) == ["A", "B", "C", "D"]
    # A has no regulator -> corrected to all four
    assert list(core.reactions["A"]) == ["A", "B", "C", "D"]
    assert list(core.reactions["B"]) == ["A", "C"]
    assert list(core.reactions["C"]) == ["A", "B"]
    assert list(core.reactions["D"]) == ["C"]
    assert core.react_count == 5 + 4
    plain = CoreSystem(small_network(), correction=False)
    assert len(plain.reactions["A"]) == 0 and plain.react_count == 5


def test_core_system_id_maps_and_str():
    core = CoreSystem(small_network())
    assert dict(core.from_tfacts_to_ids()) == {"A": 0, "B": 1, "C": 2, "D": 3}
    assert dict(core.from_ids_to_tfacts())[2] == "C"
    ids = SortedDict(core.from_reactions_to_ids())
    assert list(ids[1]) == [0, 2] and list(ids[3]) == [2]
    back = dict(core.from_reactions_ids_to_str(ids))
    assert list(back["B"]) == ["A", "C"]
    with pytest.raises(AttributeError):
        dict(core.from_reactions_ids_to_str({9: [0]}))
    text = str(core)
    assert text.startswith("== Core ==") and "B --> A, C" in text


def test_default_bounds():
    lb, ub = cf.default_bounds(3, 4)
    assert lb == [0.0] * 6 + [-20.0] * 4
    assert ub == [20.0] * 10


def test_generate_reactions_vector_layout():
    reactions = SortedDict({0: SortedSet([0, 2]), 1: SortedSet([1]), 2: SortedSet()})
    vec, mask = cf.generate_reactions_vector(reactions)
    assert vec.tolist() == [-1, 0, -1, 0, -1, 0, 0, 0, 0]
    assert mask.tolist() == [True, False, True, False, True, False, False, False, False]


def test_remove_lowest_reaction():
    reactions = SortedDict({0: SortedSet([0, 2]), 1: SortedSet([1]), 2: SortedSet([0, 1])})
    # n = 3: lambda(3) vmax(3) then k for edges (0,0) (0,2) (1,1) (2,0) (2,1)
    vres = np.array([1, 1, 1, 2, 2, 2, 5.0, -4.0, 0.5, 3.0, -0.1])
    reduced, pruned = cf.remove_lowest_reaction(vres, reactions)
    assert reduced.tolist() == [1, 1, 1, 2, 2, 2, 5.0, -4.0, 0.5, 3.0]
    assert list(pruned[2]) == [0] and list(reactions[2]) == [0, 1]
    reduced2, pruned2 = cf.remove_lowest_reaction(reduced, pruned)
    assert list(pruned2[1]) == [] and reduced2.size == 9


def write_tsv(path, header, rows):
    with open(path, "w") as fh:
        fh.write("\t".join(header) + "\n")
        for row in rows:
            fh.write("\t".join(str(v) for v in row) + "\n")


def th2_files(th2, tmp_path):
    names = [str(s) for s in th2["tfacts"]]
    net = tmp_path / "net.tsv"
    write_tsv(net, ["source", "target", "n_targets"],
              [(s, t, 1) for s, t in zip(th2["edges_source"], th2["edges_target"])])
    data_files = []
    T = th2["raw_data"].shape[2]
    for rep, mat in enumerate(th2["raw_data"]):
        f = tmp_path / ("data_p%d.csv" % rep)
        write_tsv(f, ["source"] + ["t%d" % i for i in range(T)],
                  [[names[g]] + [repr(float(v)) for v in mat[g]] for g in range(len(names))])
        data_files.append(str(f))
    tfile = tmp_path / "time.csv"
    write_tsv(tfile, ["t%d" % i for i in range(T)], [[repr(float(v)) for v in th2["times"]]])
    pfile = tmp_path / "knock.csv"
    write_tsv(pfile, names + ["t_start", "t_end"],
              [[repr(float(v)) for v in row] for row in th2["pert"]])
    return str(net), {"data": data_files, "time": str(tfile), "perturbations": str(pfile)}


def test_loaders_reproduce_reference_arrays(th2, tmp_path):
    """Files written from the golden raw arrays, parsed by this package's
    loaders, must give the arrays the reference's own loaders produced."""
    net, files = th2_files(th2, tmp_path)
    core = cc.create_core(net)
    assert list(core.tfacts) == [str(s) for s in th2["tfacts"]]
    assert core.react_count == int(th2["react_count"])
    data, sigma, times, y0 = cc.data_parsing(files)
    np.testing.assert_allclose(data, th2["data"], rtol=1e-14)
    np.testing.assert_allclose(sigma, th2["sigma"], rtol=1e-13)
    np.testing.assert_array_equal(times, th2["times"])
    np.testing.assert_allclose(y0, th2["y0"], rtol=1e-14)
    ok, pert = cc.data_parse_perturbations(files, core)
    assert ok
    np.testing.assert_allclose(pert, th2["pert"], rtol=1e-15)
    vec, mask = cf.generate_reactions_vector(SortedDict(core.from_reactions_to_ids()))
    np.testing.assert_array_equal(mask, th2["mask"])
    np.testing.assert_array_equal(vec, th2["vector_map"])
    ok, _ = cc.data_parse_perturbations({"time": files["time"]}, core)
    assert not ok


def test_loader_header_errors(tmp_path):
    bad = tmp_path / "bad.tsv"
    write_tsv(bad, ["from", "to"], [("A", "B")])
    with pytest.raises(AttributeError):
        fmt.parse_network(str(bad))
    write_tsv(bad, ["gene", "t0", "t1"], [("A", 1, 2)])
    with pytest.raises(AttributeError):
        fmt.parse_network_data(str(bad))
    write_tsv(bad, ["source", "t0", "t3"], [("A", 1, 2)])
    with pytest.raises(AttributeError):
        fmt.parse_network_data(str(bad))
    write_tsv(bad, ["t0", "x1"], [(0, 1)])
    with pytest.raises(AttributeError):
        fmt.parse_time_sequence(str(bad))
    write_tsv(bad, ["A", "B", "t_start", "stop"], [(1, 2, 0, 24)])
    with pytest.raises(AttributeError):
        fmt.parse_perturbations_data(str(bad))


def test_factory_labels():
    assert "DE" in OptimizationFactory.labels_cus()
    with pytest.raises(KeyError):
        OptimizationFactory.new_base_optimization("NOT_AN_ALGORITHM")
    opt = OptimizationFactory.new_base_optimization("PSO")
    ctx = LassimContext(None, [OptimizationArgs("PSO", {}, 1, 1, 8, 0.0)], None, None, None)
    with pytest.raises(NotImplementedError):
        opt.build(ctx, None, None, None, None, None)


def test_context_and_args():
    with pytest.raises(ValueError):
        LassimContext(None, [], None, None, None)
    args = OptimizationArgs("DE", {"gen": 2, "bogus": 1}, 4, 3, 10, 0.25)
    ctx = LassimContext("core", [args], "ode", "pert", dict)
    assert ctx.primary_first is args and ctx.secondary_first is None
    assert (ctx.ode, ctx.perturbation, ctx.SolutionClass) == ("ode", "pert", dict)
    assert (args.num_islands, args.num_evolutions, args.num_individuals) == (4, 3, 10)
    args.num_islands = 0
    assert args.num_islands >= 1
    opt = OptimizationFactory.new_base_optimization("DE")
    assert opt.handle_parameters(args) == {"gen": 2}


@pytest.mark.parametrize("gen, header", [("gen_rk_steppers", "rk_steppers.cuh"),
                                         ("gen_dop853_dense", "dop853_dense_tables.cuh")])
def test_generated_headers_are_current(gen, header, tmp_path):
    """The committed headers are what their generators write (tools/gen_*.py): nobody edited
    one by hand or changed a generator without regenerating."""
    import importlib.util
    import os
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location(gen, os.path.join(repo, "tools", gen + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = tmp_path / header
    mod.main(str(out))
    with open(os.path.join(repo, "lassim_b200", "csrc", header)) as fh:
        assert out.read_text() == fh.read()

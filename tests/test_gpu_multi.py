"""Islands sharded over two GPUs (one process per GPU, NCCL): skipped on a
one-GPU box.  The host logic is covered on CPU by the two-rank gloo test in
test_de_archipelago.py."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import REPO

pytestmark = pytest.mark.gpu

SCRIPT = r"""
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, {repo!r})
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from lassim_b200.core.core_problem import CoreProblemFactory
from lassim_b200.core.functions.common_functions import odeint1e8_lassim
from lassim_b200.core.functions.perturbation_functions import perturbation_func_sequential
from lassim_b200.pygmo import algorithm, archipelago, topology
p = dict(np.load(os.path.join({repo!r}, "tests", "golden", "th2_problem.npz")))
prob = CoreProblemFactory.new_instance((p["data"], p["sigma"], p["times"], p["pert"]), p["y0"],
                                       odeint1e8_lassim, perturbation_func_sequential, 0.5).build(
    87, (list(p["lb"]), list(p["ub"])), (p["vector_map"], p["mask"]))
archi = archipelago(algorithm.de(gen=2, variant=3, f=0.85, cr=1.0), prob, 4, 8,
                    topology=topology.ring(), seed=21)
archi.evolve(2)
archi.join()
f = np.array([c.f[0] for c in archi.champions])
assert archi.evaluations == 2 * 8 * (1 + 4), archi.evaluations     # own islands only
x0 = np.array(archi.champions[0].x)
assert prob.objfun_batch(x0[None])[0] == f[0]                       # remote champion is real
np.save(os.path.join({out!r}, "f_%d.npy" % dist.get_rank()), f)
dist.destroy_process_group()
"""


def test_islands_on_two_gpus(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "run.py"
    script.write_text(SCRIPT.format(repo=REPO, out=str(tmp_path)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    f0, f1 = np.load(tmp_path / "f_0.npy"), np.load(tmp_path / "f_1.npy")
    np.testing.assert_array_equal(f0, f1)
    assert np.isfinite(f0).all()


DEVICE_SCRIPT = r"""
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, {repo!r})
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from lassim_b200.core.core_problem import CoreProblemFactory
from lassim_b200.core.functions.common_functions import odeint1e8_lassim
from lassim_b200.core.functions.perturbation_functions import perturbation_func_sequential
from lassim_b200.pygmo import algorithm, device_archipelago, topology
p = dict(np.load(os.path.join({repo!r}, "tests", "golden", "th2_problem.npz")))
prob = CoreProblemFactory.new_instance((p["data"], p["sigma"], p["times"], p["pert"]), p["y0"],
                                       odeint1e8_lassim, perturbation_func_sequential, 0.5).build(
    87, (list(p["lb"]), list(p["ub"])), (p["vector_map"], p["mask"]))
out = {{}}
for delayed in (True, False):
    archi = device_archipelago(algorithm.de(gen=2, variant=3, f=0.85, cr=1.0), prob, 4, 8,
                               topology=topology.ring(), seed=21, delayed_migration=delayed)
    archi.evolve(3)
    archi.join()
    f = np.array([c.f[0] for c in archi.champions])
    assert archi.evaluations == 2 * 8 * (1 + 6), archi.evaluations     # own islands only
    x0 = np.array(archi.champions[0].x)
    assert prob.objfun_batch(x0[None])[0] == f[0]                       # remote champion is real
    out[delayed] = f
np.save(os.path.join({out!r}, "fd_%d.npy" % dist.get_rank()), np.stack([out[True], out[False]]))
dist.destroy_process_group()
"""


def test_device_islands_on_two_gpus(tmp_path):
    """The device archipelago across two GPUs, migration one round behind and in lock-step:
    every rank ends with the same champions, and every champion is a real evaluation."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "run.py"
    script.write_text(DEVICE_SCRIPT.format(repo=REPO, out=str(tmp_path)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29543", str(script)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    f0, f1 = np.load(tmp_path / "fd_0.npy"), np.load(tmp_path / "fd_1.npy")
    np.testing.assert_array_equal(f0, f1)
    assert np.isfinite(f0).all()

"""bench.py's CPU reference arm runs without a GPU and prints the contract's keys."""
import json
import os
import subprocess
import sys

from conftest import REPO


def test_reference_arm_json_line():
    res = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True,
                         timeout=900)
    assert res.returncode == 0, res.stderr[-2000:]
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "evals/s"
    assert line["value"] > 0 and line["higher_is_better"] is True
    cb = line["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "evals/s",
                           "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "Th2" in line["config"]["workload"]


def test_other_ranks_of_the_reference_arm_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    res = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True,
                         timeout=300, env=env)
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_parity_block_of_the_bench_line():
    """bench.py's `parity` block on synthetic 'device results': the reference's own outputs pass
    every gate; a time-series cost off by 1e-5 on a converged row does not."""
    import sys
    import numpy as np
    sys.path.insert(0, REPO)
    import bench
    g = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_uniform_full.npz")))
    E = np.concatenate([np.zeros((1024, 1, 12)), g["tight_ctrl_end"], g["tight_knock_end"]], axis=1)
    out = {"cost": g["ref_total"].copy(), "cost_main": g["ref_cost_main"].copy(),
           "pert_cost": g["ref_pert_cost"].copy(), "status": np.zeros(1024, dtype=np.int32), "E": E}
    block = bench.parity_block(out)
    assert block["ok"] and block["rows"] == 1024
    assert block["converged_rows"]["cost_main"] == 997
    assert block["knock_down_end_states"]["rows_within_1e-5_of_converged"] == 1024
    assert block["nan_rows_gpu"] == [] and block["nan_rows_agree"]
    out["cost_main"][0] *= 1 + 1e-5
    assert not bench.parity_block(out)["ok"]
    out["cost_main"][0] = g["ref_cost_main"][0]
    out["status"][5] = 1
    assert not bench.parity_block(out)["ok"]

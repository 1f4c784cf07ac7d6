#!/usr/bin/env python
"""
bench.py - cost-function evaluations per second on BASELINE.json's config 2:
the Th2 core problem (12 genes, 63 edges, 6 sample times, 12 knock-down rows ->
15 ODE solves per candidate), one DE population of 1024 candidates per GPU.

  python bench.py --gpus N --steps K --warmup W            this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    the reference's CPU path

One "step" = one pass of the hot path over one population (1024 candidates =
15,360 trajectories + the fused cost reductions).  With N > 1 (torchrun) every
rank evaluates its own copy of that population (islands are independent: weak
scaling with equal work per GPU, no data-path collective); `value` = candidates
all ranks evaluated / max-over-ranks device time.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      FP64-FMA roofline of the evaluation kernel (the path is
                compute-bound: ~0.7 KB of HBM traffic per candidate against
                ~6e7 flops).  achieved = step attempts counted by
                the kernels x algorithmic flops per attempt (lassim_flops_per_attempt
                for the explicit pair, lassim_series_flops_per_attempt for the attempts
                of the stage-parallel kernel) / device time of the evaluation; peak = DFMA
                throughput measured in this run by a register-resident probe
                kernel (MEASURED_PEAKS.json holds no FP64 figure).
  cpu_baseline  the reference's own compiled Cython functions (oracle/_ref) -
                or the oracle port when they are absent - on all host cores,
                on a bounded sample of the same population.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

POP = 1024
PERT_FACTOR = 0.5
METRIC = "cost-function evals/sec (batched GRN ODE solves)"
UNIT = "evals/s"


def load_th2():
    return dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))


def population(p, seed):
    return np.random.default_rng(seed).uniform(p["lb"], p["ub"], size=(POP, p["lb"].size))


# --------------------------------------------------------------------------
# CPU reference arm (the only place outside tests/ that executes oracle/)
# --------------------------------------------------------------------------
_W = {}


def _cpu_init(p):
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import lassim_oracle as orc
    np.seterr(all="ignore")
    _W["p"] = p
    _W["orc"] = orc
    _W["ref"] = orc.load_reference_functions()


def _cpu_eval(x):
    p, orc = _W["p"], _W["orc"]
    if _W["ref"] is not None:
        return orc.reference_total_cost(x, p["data"], p["sigma"], p["times"], p["y0"],
                                        p["pert"], PERT_FACTOR, p["vector_map"], p["mask"],
                                        _W["ref"])[0]
    return orc.total_cost(x, p["data"], p["sigma"], p["times"], p["y0"], p["pert"],
                          PERT_FACTOR, p["vector_map"], p["mask"])[0]


def cpu_kind():
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import lassim_oracle as orc
    return "reference" if orc.load_reference_functions() is not None else "port"


class CpuArm:
    """All-host-cores pool running the reference path on samples of X."""

    def __init__(self, p):
        from multiprocessing import get_context
        self.cores = os.cpu_count() or 1
        self.pool = get_context("fork").Pool(self.cores, initializer=_cpu_init, initargs=(p,))
        self.kind = cpu_kind()

    def run(self, X):
        t0 = time.perf_counter()
        costs = self.pool.map(_cpu_eval, list(X), chunksize=1)
        return time.perf_counter() - t0, np.array(costs)

    def close(self):
        self.pool.close()
        self.pool.join()


def sample_size(cores, per_core):
    return int(min(POP, max(8, per_core * cores)))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    p = load_th2()
    X = population(p, 0)
    arm = CpuArm(p)
    S = sample_size(arm.cores, 4)
    for w in range(args.warmup):
        arm.run(X[:max(arm.cores, 8)])
    total = 0.0
    for k in range(args.steps):
        lo = (k * S) % POP
        idx = (lo + np.arange(S)) % POP
        dt, _ = arm.run(X[idx])
        total += dt
    arm.close()
    value = S * args.steps / total
    sample = ("%d candidates per step (consecutive rows of the seed-0 population of %d), "
              "%s LSODA path, one worker per core" % (S, POP,
              "reference's compiled Cython functions (oracle/_ref)" if arm.kind == "reference"
              else "oracle port (scipy.integrate.odeint)"))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(1, "cpu"),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores,
                         "kind": arm.kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------
def workload_config(n_gpus, l2_note):
    return {"workload": "Th2 core (examples/Th2_lassim_core_TF_network.tsv + data_p1-p4 + "
                        "knock_data): n=12, nnz=63, dim=87, T=6, 15 ODE solves per candidate; "
                        "DE population 1024 per GPU, uniform in default bounds, default_rng(0) on every rank",
            "population_per_gpu": POP, "trajectories_per_candidate": 15,
            "pert_factor": PERT_FACTOR, "rtol": 1.49012e-8, "atol": 1.49012e-8,
            "parallelism": "islands x%d (independent populations, no collective)" % n_gpus,
            "l2": l2_note}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks line sampled while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "reasons": sorted(reasons), "samples": len(sm)}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from lassim_b200 import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback "
                         "(use --impl reference for the CPU reference arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    p = load_th2()
    X = population(p, 0)      # the config-2 population on every rank: equal work per GPU
    eng = Engine(local)
    eng.set_method(args.method)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], PERT_FACTOR)
    flops_per_attempt = eng.flops_per_attempt      # SURVEY.md 8d, per integrator
    # long time series leave the standard kernel after a cap of attempts and continue in the
    # stage-parallel Gauss kernel (two warps per trajectory), with its own attempts and flops
    series_flops = eng.series_flops_per_attempt(POP)

    dX = torch.from_numpy(X).cuda()
    dcost = torch.empty(POP, dtype=torch.float64, device="cuda")
    dcnt = torch.zeros(POP * 3, dtype=torch.int64, device="cuda")
    dstatus = torch.zeros(POP, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    # a side stream: the legacy default stream's handle is 0, which the C ABI
    # reads as "the engine's own stream"; events must be on the launch stream
    side = torch.cuda.Stream()
    torch.cuda.set_stream(side)
    stream = side.cuda_stream
    assert stream != 0

    def step_device():
        eng.eval_device(POP, dX.data_ptr(), dcost.data_ptr(), None, None,
                        dstatus.data_ptr(), dcnt.data_ptr(), stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    fp64_peak = eng.fp64_peak_tflops(400000) if rank == 0 else None

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step_device()
    torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = eng.launch_count
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    kernel_ms, attempts, sp_ms, sp_att = [], [], [], []
    barrier()
    for k in range(args.steps):
        flush.zero_()                      # evict L2 between timed steps
        starts[k].record()
        step_device()
        ends[k].record()
        ends[k].synchronize()
        kernel_ms.append(eng.last_kernel_ms())
        attempts.append(int(dcnt.view(POP, 3)[:, :2].sum().item()))
        sp_ms.append(eng.last_stage_parallel_ms())            # the dominant kernel's own launch
        sp_att.append(eng.last_stage_parallel_attempts())
    barrier()
    launches = eng.launch_count - launches0
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    # A knock-down cost is NaN in the reference too when a control end state's rounding-level
    # remnant (1e-190) is negative under a knocked state of 1e120 (-inf/inf after the clip,
    # perturbation_functions.pyx:86-95): every integration must succeed, all but such rows be finite
    nonfinite = int((~torch.isfinite(dcost)).sum().item())
    ok = bool((dstatus == 0).all().item()) and nonfinite <= POP // 200

    # the other Dormand-Prince pair on the same population (outside the timed region):
    # the headline uses --method (default 8(5,3)); BASELINE.json's north_star names 5(4)
    other = None
    if rank == 0 and world == 1:
        alt = "dp45" if args.method == "dop853" else "dop853"
        eng.set_method(alt)
        ms_alt = []
        for _ in range(3):
            step_device()
            torch.cuda.synchronize()
            ms_alt.append(eng.last_kernel_ms())
        other = {"integrator": {"dop853": "Dormand-Prince 8(5,3)", "dp45": "Dormand-Prince 5(4)"}[alt],
                 "kernel_ms": float(np.median(ms_alt[1:])),
                 "evals_per_s": POP / (float(np.median(ms_alt[1:])) * 1e-3)}
        eng.set_method(args.method)

    # the same kernel with the machine kept full (16 copies of the population):
    # separates kernel quality from the single-population critical path
    sat = None
    if rank == 0:
        reps = 16
        dXs = dX.repeat(reps, 1).contiguous()
        dcs = torch.empty(POP * reps, dtype=torch.float64, device="cuda")
        dns = torch.zeros(POP * reps * 3, dtype=torch.int64, device="cuda")
        for _ in range(2):
            eng.eval_device(POP * reps, dXs.data_ptr(), dcs.data_ptr(), None, None, None,
                            dns.data_ptr(), stream)
        torch.cuda.synchronize()
        ms_sat = eng.last_kernel_ms()
        att = int(dns.view(-1, 3)[:, :2].sum().item())
        att_series = eng.last_stage_parallel_attempts()
        sat_flops = (att - att_series) * flops_per_attempt + \
            att_series * eng.series_flops_per_attempt(POP * reps)
        sat = {"candidates": POP * reps, "kernel_ms": ms_sat,
               "evals_per_s": POP * reps / (ms_sat * 1e-3),
               "time_series_kernel": eng.series_kernel_name(POP * reps),
               "achieved_tflops": sat_flops / (ms_sat * 1e-3) / 1e12}
        del dXs, dcs, dns

    # end to end through the public host API (C-ABI lassim_eval, host buffers)
    Xpin = torch.from_numpy(X).pin_memory()
    Xh = Xpin.numpy()
    for _ in range(2):
        eng.eval(Xh)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        cost_h = eng.eval(Xh)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([dev_ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_ms_max = float(t[0]), float(t[1])

    if rank == 0:
        value = world * POP * args.steps / (dev_ms_max * 1e-3)
        e2e_value = world * POP * args.steps / (e2e_ms_max * 1e-3)
        k_ms = float(np.mean(kernel_ms))
        series_attempts = float(np.mean(sp_att))
        flops = (float(np.mean(attempts)) - series_attempts) * flops_per_attempt + \
            series_attempts * series_flops
        achieved_all = flops / (k_ms * 1e-3) / 1e12
        # the dominant kernel: the stage-parallel kernel the long time series are handed to
        d_ms = float(np.mean(sp_ms))
        achieved = series_attempts * series_flops / (d_ms * 1e-3) / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": dict(workload_config(world, "flushed between timed steps (256 MiB write); "
                                                  "per-step CUDA events on the launching stream"),
                           integrator={"dop853": "Dormand-Prince 8(5,3)",
                                       "dp45": "Dormand-Prince 5(4)"}[args.method]),
            "results_ok": ok, "nonfinite_costs": nonfinite,
            "e2e": {"value": e2e_value, "unit": UNIT,
                    "h2d_bytes_per_step": int(X.nbytes), "d2h_bytes_per_step": int(POP * 8),
                    "api": "lassim_eval (C ABI) with pinned host X, host cost out"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "fp64", "achieved": achieved, "peak": fp64_peak,
                         "unit": "TFLOP/s", "frac": achieved / fp64_peak if fp64_peak else None,
                         # dram__bytes_read.sum + dram__bytes_write.sum of the evaluation's three
                         # kernels, one ncu --set full capture (profiles/r1_m_ncu_full_summary.txt);
                         # algorithmic bytes are X in + cost out = 0.72 MB: the path is compute-bound
                         "traffic": 514816,
                         "traffic_source": "profiles/r1_m_ncu_full_summary.txt (this kernel; the "
                                           "evaluation's three kernels together: 2,266,368 bytes)",
                         "kernel": eng.series_kernel_name(POP), "kernel_ms": d_ms,
                         "attempts_per_launch": series_attempts,
                         "flops_per_attempt": series_flops,
                         "share_of_step": d_ms / k_ms,
                         "note": "dominant kernel of the step: 2-3 limit-cycle time series of 12,000 "
                                 "sequential attempts each set its duration (latency-bound)",
                         # all kernels of one evaluation together (three launches, two streams)
                         "evaluation": {"kernels": [eng.kernel_name + " (time series to the cap)",
                                                    eng.series_kernel_name(POP),
                                                    eng.kernel_name + " (knock-downs, second stream)"],
                                        "ms": k_ms, "achieved": achieved_all,
                                        "frac": achieved_all / fp64_peak if fp64_peak else None,
                                        "step_attempts": float(np.mean(attempts)),
                                        "explicit_pair_flops_per_attempt": flops_per_attempt},
                         "peak_source": "in-run DFMA probe kernel (MEASURED_PEAKS.json has "
                                        "HBM and bf16 figures only)",
                         "machine_full": dict(sat, frac=sat["achieved_tflops"] / fp64_peak,
                                              note="same kernel on 16 copies of the population "
                                                   "(outside the timed region): what the kernel "
                                                   "does when the batch outlasts its longest "
                                                   "trajectory")},
        }
        if other is not None:
            line["other_integrator"] = other
        if world == 1 and not args.no_cpu_baseline:
            arm = CpuArm(p)
            S = sample_size(arm.cores, 32)   # about 10 s on 16 cores: includes the slow candidates
            dt, costs = arm.run(X[:S])
            arm.close()
            gpu = cost_h[:S]
            line["cpu_baseline"] = {
                "value": S / dt, "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
                "sample": "first %d candidates of the same population, one pass, one worker "
                          "per core (%.1f s)" % (S, dt)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--method", default=os.environ.get("LASSIM_METHOD", "dop853"),
                    choices=["dop853", "dp45"], help="integrator of the CUDA path")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Benchmark of the OptiMo objective hot path on B200 (contract: see the task statement, section 4).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload stick240|tree64|grad]

A "step" is one pass of the hot path over one batch of synthetic candidates:
  workload stick240 (default; BASELINE.json configs[3], the one `north_star`'s target is quoted on):
      4096 candidate edits of the stick-figure skeleton x 240 frames  -> 4096 x 1216 = 4.98 M sample-evals / GPU / step
  workload grad : the 937-column finite-difference batch of one objective+gradient evaluation of that scene
  workload grad_stick / grad_conductor: the same for the shipped stick-figure / conductor scenes (configs[0] / configs[2])
  workload tree64: BASELINE.json configs[4], 64-joint tree x 2000 frames, 256 candidates
N > 1 (torchrun, one process per GPU): every rank evaluates its own candidates (weak scaling, no data-path collective);
the per-candidate scalars are all-gathered over NCCL inside the timed step.

The line printed by rank 0 carries `value` (device-resident inputs, CUDA-event timed), `e2e` (through
Core.CalculateFullObjectiveBatch with pinned HOST buffers), `roofline` (FP64 pipe, against the DFMA rate measured
in this run) and `cpu_baseline` (the CPU oracle timed on this box's host cores on a bounded sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "rnea_sample_evals_per_sec"
UNIT = "sample-evals/s"


def load_stick():
    with open(os.path.join(ROOT, "tests", "golden", "scenes", "stick-figure.json")) as fh:
        return json.load(fh)


def make_workload(name, rank=0):
    from optimo_b200 import synthetic
    if name in ("stick240", "grad"):
        scene = synthetic.stick_240(load_stick())
        B = 4096
        desc = "4096 stick-figure candidate edits x 240 frames (BASELINE.json configs[3])"
    elif name == "tree64":
        scene = synthetic.tree_scene(64, 2000)
        B = 256
        desc = "64-joint tree x 2000 frames, 256 candidates (BASELINE.json configs[4])"
    elif name in ("grad_stick", "grad_conductor"):
        # BASELINE.json configs[0] / configs[2]: one objective + finite-difference gradient of a shipped scene
        with open(os.path.join(ROOT, "tests", "golden", "scenes", {"grad_stick": "stick-figure", "grad_conductor": "conductor"}[name] + ".json")) as fh:
            scene = json.load(fh)
        B = 1
        desc = {"grad_stick": "resources/scene-stick-figure.json (BASELINE.json configs[0])",
                "grad_conductor": "resources/scene-conductor.json, IK handle + 2 Gaussian kernels + w_traj 12.5 (BASELINE.json configs[2])"}[name]
    else:
        raise SystemExit(f"unknown workload {name}")
    return scene, B, desc


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = threading.Event()

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
            "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for k, n in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(self.rows)}


def cpu_reference_leg(scene, B_sample, threads, x0, first=0):
    """The reference's CPU implementation of the path: oracle/_ref (the reference's own RBDL / curve objects) when it
    was built, else the restatement. Candidate-parallel over all host threads (figure ii)."""
    from optimo_b200 import synthetic
    from oracle import oracle as O
    kind = "ref" if O.available("ref") else "port"
    if kind == "port" and not O.available("port"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
    o = O.Oracle(scene, kind)
    X = synthetic.perturbed_candidates(x0, B_sample, first=first)
    cores = O.hardware_concurrency(kind) if threads == 0 else threads
    sec, f = o.bench_batch(X, 1, cores)
    n_s = O.time_grid(*o.window, kind=kind).size
    # BASELINE.md section 3, figure i ("ref-as-is"): what OptiMo does today -- candidates one after another, parallel_for over the
    # time samples for the curve / FK sampling (core_rbdl.cpp:158), serial RNEA loop (core_rbdl.cpp:338). Small sample: it is slow.
    n_as_is = max(2, min(B_sample, 16))
    sec0, f0 = o.bench_batch(X[:n_as_is], 0, cores)
    assert np.array_equal(f0, f[:n_as_is]) or np.allclose(f0, f[:n_as_is], rtol=1e-12, atol=0)
    return {"kind": "reference" if kind == "ref" else "port", "cores": cores, "seconds": sec, "candidates": B_sample, "n_s": int(n_s), "f": f, "X": X,
            "as_is": {"value": n_as_is * int(n_s) / sec0, "unit": UNIT, "objective_evals_per_s": n_as_is / sec0,
                      "sample": f"{n_as_is} candidates one after another, parallel_for over samples on {cores} threads, serial RNEA (the reference's own schedule)"}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from optimo_b200.core import Core
    scene, B, desc = make_workload(args.workload)
    x0 = Core(scene).GetDescriptor()         # host-only use of Core: descriptor packing, no CUDA
    from oracle import oracle as O
    cores = O.hardware_concurrency("port") if O.available("port") else os.cpu_count()
    sample = max(cores, min(B, 64 * cores))
    times = []
    leg = None
    for it in range(args.warmup + args.steps):
        leg = cpu_reference_leg(scene, sample, cores, x0)
        if it >= args.warmup:
            times.append(leg["seconds"])
    sec = float(np.mean(times))
    value = sample * leg["n_s"] / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "candidates_per_step": sample, "samples_per_candidate": leg["n_s"]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": leg["cores"], "kind": leg["kind"],
                         "sample": f"{sample} of {B} candidates per step, candidate-parallel over {leg['cores']} host threads",
                         "as_is": leg["as_is"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "objective_evals_per_s": sample / sec, "frame_evals_per_s": value / 5.0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist

    from optimo_b200 import synthetic
    from optimo_b200.core import Core

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene, B, desc = make_workload(args.workload, rank)
    core = Core(scene, device=local)
    if args.workload == "grad_conductor":
        core.set_kernels([(1.0, 20.0, 2.6), (1.0, 10.0, 1.0)])
        core.UpdateGlobalWeight(12.5)
    x0 = core.GetDescriptor()
    lo, hi = core.window
    eng = core.engine(synced_window=(lo, hi))
    N_s, D, J = eng.num_samples, eng.D, core.J
    idx = core.free_indices()

    grad = args.workload.startswith("grad")
    if grad:
        B = 2 * idx.size + 1
        x_base = synthetic.perturbed_candidates(x0, 1, first=0)[0]
        X = np.tile(x_base, (B, 1))
        for k, i in enumerate(idx):
            X[2 * k + 1, i] += 1e-4
            X[2 * k + 2, i] -= 1e-4
        desc = (f"one objective+gradient evaluation = {B} finite-difference columns of "
                + ("the stick-figure x 240 frames scene" if args.workload == "grad" else desc) +
                ", through omo_eval_grad_columns (columns evaluated only on the tiles their control point reaches; value = executed sample-evals/s)")
    else:
        X = synthetic.perturbed_candidates(x0, B, first=rank * B)

    stream = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local))
    dX = torch.empty((B, D), dtype=torch.float64, device="cuda")
    dF = torch.empty(B, dtype=torch.float64, device="cuda")
    dAll = torch.empty(B * world, dtype=torch.float64, device="cuda") if world > 1 else None
    flush = torch.empty(256 * 1024 * 1024 // 8, dtype=torch.float64, device="cuda")   # 256 MiB > 126 MB L2
    Xp = eng.pinned_array((B, D))
    Xp[:] = X
    Fp = eng.pinned_array((B,))
    eng.h2d(dX.data_ptr(), X)

    def step_device():
        if grad:   # the finite-difference columns through the API: built on the device, sparse over tiles (include/omo.h)
            eng.eval_grad_columns(x_base, idx, 1e-4, 0, B, dF.data_ptr(), sync=False)
        else:
            eng.eval_device(dX.data_ptr(), B, dF.data_ptr(), None, sync=False)
        if world > 1:
            dist.all_gather_into_tensor(dAll, dF)

    def step_e2e():
        if grad:   # host x in, (f0, gradient) out on the host
            g0 = eng.eval_grad(x_base, idx, 1e-4)
            Fp[0] = g0[0]; Fp[1::2] = g0[2][:, 0]; Fp[2::2] = g0[2][:, 1]
        else:
            eng.eval_host_ptr(Xp.ctypes.data, B, Fp.ctypes.data)     # H2D of X, kernels, D2H of f, all inside
        if world > 1:
            dist.barrier()

    def timed(fn, steps, warmup, kernel_timing=False):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                fn()
            stream.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            if kernel_timing:
                eng.kernel_timing(reset=1)
            l0 = eng.counters()["launches"]
            evs = []
            t0 = time.perf_counter()
            for _ in range(steps):
                flush.zero_()                                     # evict L2 between timed iterations (not timed)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                fn()
                e1.record(stream)
                evs.append((e0, e1))
            stream.synchronize()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            wall = time.perf_counter() - t0
            ms = sum(a.elapsed_time(b) for a, b in evs) / steps
            kt = eng.kernel_timing(reset=-1) if kernel_timing else (0.0, 0)
            launches = eng.counters()["launches"] - l0
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, wall, kt, launches

    Xfp = Fp2 = None
    if not grad:   # the same candidates as their free entries only (omo_eval_free: expansion over x0 on the device)
        Xfp = eng.pinned_array((B, idx.size))
        Xfp[:] = X[:, idx]
        Fp2 = eng.pinned_array((B,))

    def step_e2e_free():
        eng.eval_free_host_ptr(x0, idx, Xfp.ctypes.data, B, Fp2.ctypes.data)
        if world > 1:
            dist.barrier()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_dev, wall_dev, (k_ms, k_n), launches = timed(step_device, args.steps, args.warmup, kernel_timing=True)
    ms_e2e, wall_e2e, _, _ = timed(step_e2e, args.steps, max(args.warmup, 3))
    ms_free = None
    if not grad:
        ms_free, _, _, _ = timed(step_e2e_free, args.steps, max(args.warmup, 3))
    if sampler:
        sampler.stop_flag.set()
        sampler.join(timeout=2)

    # correctness of what was timed: device-resident and host paths agree bit for bit
    f_dev = dF.cpu().numpy()
    assert np.array_equal(f_dev, np.asarray(Fp)), "device-resident and host-buffer paths disagree"
    if Fp2 is not None:
        assert np.array_equal(f_dev, np.asarray(Fp2)), "free-parameter and full-descriptor host paths disagree"

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # executed curve work -> ALGORITHMIC FLOPs per sample-eval
    eng.set_counting(True)
    eng.eval_device(dX.data_ptr(), B, dF.data_ptr(), None, sync=True)
    cnt = eng.counters()
    eng.set_counting(False)
    n_se = B * N_s
    executed_frac = 1.0
    if grad:   # SURVEY.md section 8(d) dense counting rule: report the sample-evals actually executed
        ex_items, dense_items = eng.grad_sparsity()
        executed_frac = ex_items / max(dense_items, 1)
    bez, newt = cnt["bezier_evals"] / n_se, cnt["newton_iters"] / n_se
    # SURVEY.md section 8(d) formula. `flops_reference_work` counts all J-1 bodies (what the reference executes per
    # sample); `flops_executed` counts only the bodies the batch kernel walks (candidate-independent subtrees are
    # evaluated once per window, SURVEY.md Appendix B: "must then report executed FLOPs") -- `frac` uses the latter.
    walked, walked_general = eng.walked_bodies()
    flops_ref = synthetic.flops_per_sample_eval(J, bez, newt)
    flops_se = synthetic.flops_per_sample_eval(walked + 1, bez, newt)
    dfma_peak = eng.dfma_peak()                                  # measured in this run: MEASURED_PEAKS.json has no FP64 figure
    k_ms_per_launch = k_ms / max(k_n, 1)
    achieved = n_se * executed_frac * flops_se / (k_ms_per_launch * 1e-3)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peaks = json.load(fh)
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    n_tiles = (N_s + 123) // 124

    def keyed(node, root=True):
        own = 0 if root else sum(1 for k in ("rot_x", "rot_y", "rot_z") if node[k].get("curve"))
        return own + sum(keyed(ch, False) for ch in node.get("children", []))
    n_keyed = keyed(scene["joints"])
    # what the objective kernel must move: the prepared segment records (10 doubles per control point) and the tile table in,
    # the tile partials out. (X itself, 8 D bytes per candidate, is read by prepare_kernel.)
    alg_bytes = float(B) * (80.0 * (D // 5) + 2.0 * n_tiles * n_keyed + 16.0 * n_tiles)
    traffic = None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the committed `ncu --set full` capture
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            traffic = json.load(fh).get(args.workload, {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {
        "bound": "fp64", "achieved": achieved / 1e12, "peak": dfma_peak / 1e12, "unit": "TFLOP/s", "frac": achieved / dfma_peak,
        "traffic": traffic, "kernel": "omo::objective_kernel", "kernel_ms": k_ms_per_launch, "kernel_share_of_step": k_ms_per_launch / ms_dev,
        "flops_per_sample_eval": flops_se, "flops_per_sample_eval_reference_work": flops_ref,
        "frac_of_reference_work": n_se * flops_ref / (k_ms_per_launch * 1e-3) / dfma_peak,
        "bodies": {"scene": J - 1, "walked": walked, "walked_general": walked_general},
        "bezier_evals_per_sample_eval": bez, "newton_iters_per_sample_eval": newt,
        "peak_source": "dependent-free DFMA microbenchmark (omo_measure_dfma_peak) in this run",
        "hbm": {"algorithmic_bytes": alg_bytes, "algorithmic_GBps": alg_bytes / (k_ms_per_launch * 1e-3) / 1e9, "peak_GBps": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
    }

    # CPU baseline on this box's host cores, bounded sample of the same workload
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    sample = max(cores, min(B, 64 * cores))
    if grad or world > 1 or args.no_cpu_baseline:
        cpu = None
    else:
        leg = cpu_reference_leg(scene, sample, 0, x0)
        cpu_val = leg["candidates"] * leg["n_s"] / leg["seconds"]
        # parity of the timed batch against the checker on that sample
        rel = np.max(np.abs(f_dev[:sample] - leg["f"]) / np.abs(leg["f"]))
        assert rel <= 1e-9, f"GPU batch deviates from the CPU oracle: {rel}"
        cpu = {"value": cpu_val, "unit": UNIT, "cores": leg["cores"], "kind": leg["kind"],
               "sample": f"first {sample} of {B} candidates, candidate-parallel over {leg['cores']} host threads, {leg['seconds']:.2f} s",
               "max_rel_diff_vs_gpu": float(rel), "as_is": leg["as_is"]}

    n_total = n_se * executed_frac * world
    value = n_total / (ms_dev * 1e-3)
    e2e_value = n_total / (ms_e2e * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "candidates_per_gpu": B, "samples_per_candidate": N_s, "joints": J, "descriptor": D,
                   "l2": "flushed between timed iterations (256 MiB memset)", "gather": "nccl all_gather of f" if world > 1 else "none"},
        "clocks": sampler.summary() if sampler else None,
        "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": int((8 * D + 4 * idx.size) if grad else B * D * 8), "d2h_bytes_per_step": int(B * 8)},
        "e2e_free_parameters": None if ms_free is None else {
            "value": n_total / (ms_free * 1e-3), "unit": UNIT, "ms_per_step": ms_free, "api": "omo_eval_free (candidates as their F free entries, expanded on the device)",
            "h2d_bytes_per_step": int(B * idx.size * 8 + D * 8 + idx.size * 4), "d2h_bytes_per_step": int(B * 8)},
        "gpu_launches": int(launches),
        "roofline": roofline, "cpu_baseline": cpu,
        "executed_fraction_of_dense_sample_evals": executed_frac,
        "objective_evals_per_s": B * world / (ms_dev * 1e-3), "frame_evals_per_s": value / 5.0,
        "objective_plus_gradient_evals_per_s": (world / (ms_dev * 1e-3)) if grad else (B * world / (2 * idx.size + 1)) / (ms_dev * 1e-3),
        "wall_s": {"device_loop": wall_dev, "e2e_loop": wall_e2e},
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _reserve_stdout():
    """The contract is ONE JSON line on stdout. Libraries write there too (NCCL prints its version line on fd 1 when the first
    communicator comes up), so fd 1 is pointed at stderr for the run and the line goes to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(saved, "w")


def main():
    global print
    out = _reserve_stdout()
    _print = print

    def print(*a, **k):   # noqa: A001 -- only the result lines use print() in this file
        k.setdefault("file", out)
        _print(*a, **k)
        out.flush()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="stick240", choices=["stick240", "tree64", "grad", "grad_stick", "grad_conductor"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="development runs: skip the timed CPU leg (cpu_baseline = null)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

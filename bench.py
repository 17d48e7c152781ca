#!/usr/bin/env python
"""Benchmark of the OptiMo objective hot path on B200 (contract: see the task statement, section 4).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--no-workloads] [--sustain S]

A "step" is one pass of the hot path over one batch of synthetic candidates. Workloads (BASELINE.json configs):
  stick240 (default; configs[3], the one `north_star`'s target is quoted on): 4096 candidate edits of the stick-figure
      skeleton x 240 frames -> 4096 x 1216 = 4.98 M sample-evals per step
  grad           : configs[3]'s gradient variant, the 937 finite-difference columns of one objective+gradient evaluation
  grad_stick     : configs[0], one objective + gradient of resources/scene-stick-figure.json (157 columns x 215 samples)
  foxtail_solve  : configs[1], the full NLopt solve of resources/scene-fox-tail.json with the GPU callback
  grad_conductor : configs[2], one objective + gradient of resources/scene-conductor.json (IK handle, 2 kernels, w_traj 12.5)
  tree64         : configs[4], 64-joint tree x 2000 frames, 256 candidates

N = 1: the line's `value` / `e2e` / `roofline` / `cpu_baseline` are the stick240 batch; `workloads` carries one entry per
BASELINE config with value, ms_per_step, roofline.frac and max_rel_diff_vs_oracle (parity of what was timed).
N > 1 (torchrun, one process per GPU): STRONG scaling -- the ONE 4096-candidate batch is split over the ranks by
omo_shard_range, every rank evaluates its slice, f is all-gathered over NCCL inside the timed step; `weak` (4096 per rank)
and `grad_sharded` (the 937 columns split over the ranks, gathered, compared bit for bit with the one-GPU gradient) are
reported beside it, and `single_process` times the same split through the product's own multi-device entry point
(omo_create_multi: one host thread, N GPUs).
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "rnea_sample_evals_per_sec"
UNIT = "sample-evals/s"
H = 1e-4   # central-difference step


def _synthetic():
    """optimo_b200/synthetic.py loaded by path: the reference arm must not import the product package (it would map the
    product's shared libraries into the process that times the CPU reference)."""
    spec = importlib.util.spec_from_file_location("omo_synthetic", os.path.join(ROOT, "optimo_b200", "synthetic.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_scene(name):
    with open(os.path.join(ROOT, "tests", "golden", "scenes", name + ".json")) as fh:
        return json.load(fh)


CONDUCTOR_KERNELS = [(1.0, 20.0, 2.6), (1.0, 10.0, 1.0)]
W_TRAJ = {"grad_conductor": 12.5}

WORKLOADS = {
    "stick240": ("4096 stick-figure candidate edits x 240 frames (BASELINE.json configs[3])", 4096),
    "grad": ("one objective+gradient evaluation of the stick-figure x 240 frames scene = 937 finite-difference columns (BASELINE.json configs[3], gradient variant)", 1),
    "grad_stick": ("one objective+gradient evaluation of resources/scene-stick-figure.json (BASELINE.json configs[0])", 1),
    "grad_conductor": ("one objective+gradient evaluation of resources/scene-conductor.json, IK handle + 2 Gaussian kernels + w_traj 12.5 (BASELINE.json configs[2])", 1),
    "tree64": ("64-joint tree x 2000 frames, 256 candidates (BASELINE.json configs[4])", 256),
    "foxtail_solve": ("resources/scene-fox-tail.json full NLopt solve (LN_BOBYQA, 2000 evaluations) with the GPU objective callback (BASELINE.json configs[1])", 1),
}


def make_scene(name):
    syn = _synthetic()
    if name in ("stick240", "grad"):
        return syn.stick_240(load_scene("stick-figure"))
    if name == "tree64":
        return syn.tree_scene(64, 2000)
    return load_scene({"grad_stick": "stick-figure", "grad_conductor": "conductor", "foxtail_solve": "fox-tail"}[name])


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

    def stop(self):
        self.stop_flag.set()
        self.join(timeout=2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for k, n in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": sorted(reasons), "samples": len(self.rows)}


# ---------------------------------------------------------------------------------------------------------------------
# CPU side: the reference's implementation of the path (oracle/_ref = the reference's own RBDL / curve objects, else the port)
# ---------------------------------------------------------------------------------------------------------------------
def make_oracle(name, scene):
    from oracle import oracle as O
    kind = "ref" if O.available("ref") else "port"
    if kind == "port" and not O.available("port"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
    o = O.Oracle(scene, kind)
    if name == "grad_conductor":
        o.set_kernels(CONDUCTOR_KERNELS)
    o.set_weight_trajectory(W_TRAJ.get(name, 0.0))
    return O, o, kind


def cpu_leg(name, scene, X, as_is=True):
    """Candidate-parallel evaluation of the rows of X over all host threads (figure ii), plus figure i
    ("ref-as-is": candidates one after another, parallel_for over samples, serial RNEA -- what OptiMo does) on a few rows."""
    O, o, kind = make_oracle(name, scene)
    cores = O.hardware_concurrency(kind)
    sec, f = o.bench_batch(X, 1, cores)
    n_s = int(O.time_grid(*o.window, kind=kind).size)
    out = {"kind": "reference" if kind == "ref" else "port", "cores": cores, "seconds": sec, "candidates": int(X.shape[0]), "n_s": n_s, "f": f}
    if as_is:
        n_as_is = max(2, min(X.shape[0], 16))
        sec0, f0 = o.bench_batch(X[:n_as_is], 0, cores)
        assert np.allclose(f0, f[:n_as_is], rtol=1e-12, atol=0)
        out["as_is"] = {"value": n_as_is * n_s / sec0, "unit": UNIT, "objective_evals_per_s": n_as_is / sec0,
                        "sample": f"{n_as_is} candidates one after another, parallel_for over samples on {cores} threads, serial RNEA (the reference's own schedule)"}
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores. Nothing of the product
    is imported here (descriptor from the oracle, synthetic scenes loaded by file path)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload if args.workload != "foxtail_solve" else "stick240"
    syn = _synthetic()
    scene = make_scene(name)
    desc, B = WORKLOADS[name]
    O, o, kind = make_oracle(name, scene)
    x0 = o.descriptor()
    cores = O.hardware_concurrency(kind)
    if name.startswith("grad"):
        idx = o.free_indices()
        x_base = syn.perturbed_candidates(x0, 1, first=0)[0]
        X = np.tile(x_base, (2 * idx.size + 1, 1))
        for k, i in enumerate(idx):
            X[2 * k + 1, i] += H
            X[2 * k + 2, i] -= H
        sample_note = f"all {X.shape[0]} finite-difference columns"
    elif name == "tree64":
        X = syn.perturbed_candidates(x0, max(cores, 16))
        sample_note = f"first {X.shape[0]} of {B} candidates per step (one objective of this size takes a core ~2 s)"
    else:
        X = syn.perturbed_candidates(x0, B)
        sample_note = f"all {B} candidates per step"
    n_s = int(O.time_grid(*o.window, kind=kind).size)
    times = []
    for it in range(args.warmup + args.steps):
        sec, _ = o.bench_batch(X, 1, cores)
        if it >= args.warmup:
            times.append(sec)
    sec = float(np.mean(times))
    n_as_is = max(2, min(X.shape[0], 16))
    sec0, _ = o.bench_batch(X[:n_as_is], 0, cores)
    value = X.shape[0] * n_s / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": desc, "candidates_per_step": int(X.shape[0]), "samples_per_candidate": n_s},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference" if kind == "ref" else "port",
                         "sample": f"{sample_note}, candidate-parallel over {cores} host threads (parallelutil::parallel_for over candidates, a private object graph + RBD::Model per thread); RBDL on its SimpleMath backend (Eigen is not installed)",
                         "as_is": {"value": n_as_is * n_s / sec0, "unit": UNIT, "objective_evals_per_s": n_as_is / sec0,
                                   "sample": f"{n_as_is} candidates one after another, parallel_for over samples, serial RNEA (the reference's own schedule)"}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "objective_evals_per_s": X.shape[0] / sec, "frame_evals_per_s": value / 5.0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------------------------
class Gpu:
    """torch is used for the NCCL plumbing, the L2 flush buffer and CUDA events on the engine's stream; nothing else."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
        torch.cuda.set_device(self.local)
        self.cpu_group = None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.cpu_group = dist.new_group(backend="gloo")   # host-side barrier that leaves the GPUs idle
        self.flush = torch.empty(256 * 1024 * 1024 // 8, dtype=torch.float64, device="cuda")   # 256 MiB > 126 MB L2

    def stream_of(self, eng):
        return self.torch.cuda.ExternalStream(eng.stream(), device=self.torch.device("cuda", self.local))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def timed(self, eng, fn, steps, warmup, kernel_timing=False, flush=True):
        """W warm-up calls, then `steps` calls each bracketed by CUDA events on the engine's stream (L2 flushed before each,
        outside the events); barrier + synchronize on both sides; MAX over ranks. Returns (ms per step, wall s, (kernel ms,
        kernel launches), launches counted by the engine)."""
        torch = self.torch
        stream = self.stream_of(eng)
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                fn()
            stream.synchronize()
            self.barrier()
            torch.cuda.synchronize()
            if kernel_timing:
                eng.kernel_timing(reset=1)
            l0 = eng.counters()["launches"]
            evs = []
            t0 = time.perf_counter()
            for _ in range(steps):
                if flush:
                    self.flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                fn()
                e1.record(stream)
                evs.append((e0, e1))
            stream.synchronize()
            torch.cuda.synchronize()
            self.barrier()
            wall = time.perf_counter() - t0
            ms = sum(a.elapsed_time(b) for a, b in evs) / steps
            kt = eng.kernel_timing(reset=-1) if kernel_timing else (0.0, 0)
            launches = eng.counters()["launches"] - l0
        if self.world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, wall, kt, launches


_KEEP = []   # engines outlive every torch tensor that was used on their streams (torch records an event on that stream when it frees one)


def build(name, device):
    from optimo_b200.core import Core
    scene = make_scene(name)
    desc, B = WORKLOADS[name]
    core = Core(scene, device=device)
    _KEEP.append(core)
    if name == "grad_conductor":
        core.set_kernels(CONDUCTOR_KERNELS)
    core.UpdateGlobalWeight(W_TRAJ.get(name, 0.0))
    x0 = core.GetDescriptor()
    lo, hi = core.window
    eng = core.engine(synced_window=(lo, hi))
    return SimpleNamespace(name=name, scene=scene, desc=desc, B=B, core=core, eng=eng, x0=x0, lo=lo, hi=hi, N_s=eng.num_samples, D=eng.D, J=core.J,
                           idx=core.free_indices(), grad=name.startswith("grad"), fk=(W_TRAJ.get(name, 0.0) != 0.0 or name == "grad_conductor"))


def fd_columns(x_base, idx):
    X = np.tile(x_base, (2 * idx.size + 1, 1))
    for k, i in enumerate(idx):
        X[2 * k + 1, i] += H
        X[2 * k + 2, i] -= H
    return X


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh)
    except Exception:
        return {}


def ncu_capture(name):
    """dram bytes and executed FP64 instruction counts per launch of the objective kernel, from the committed `ncu --set full`
    capture of this workload (profiles/traffic.json names the report each figure was read from)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            return json.load(fh).get(name, {})
    except Exception:
        return {}


def measure_analytic_tree64(G):
    """The gradient a finite-difference batch cannot deliver: BASELINE configs[4]'s scene has 113 967 free parameters (227 935 perturbed
    descriptors of 189 945 doubles = 346 GB). omo_eval_grad_analytic (reverse mode: one forward sweep, one fold-back and their adjoints
    per sample), host x in, gradient out, wall clock; checked against central differences of 24 parameters evaluated as a batch."""
    W = build("tree64", G.local)
    core = W.core
    x = _synthetic().perturbed_candidates(W.x0, 1, first=1)[0]
    idx = W.idx
    f, g, used = core.CalculateAnalyticGradient(W.lo, W.hi, x, W.x0)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        f, g, used = core.CalculateAnalyticGradient(W.lo, W.hi, x, W.x0)
        ts.append(time.perf_counter() - t0)
    sel = np.arange(0, idx.size, max(1, idx.size // 24))[:24]
    X = np.tile(x, (2 * sel.size, 1))
    for k, sidx in enumerate(sel):
        X[2 * k, idx[sidx]] += H
        X[2 * k + 1, idx[sidx]] -= H
    fx = core.CalculateFullObjectiveBatch(W.lo, W.hi, X, W.x0)
    g_fd = (fx[0::2] - fx[1::2]) / (2 * H)
    ms = float(np.median(ts) * 1e3)
    return {"workload": "analytic gradient of the 64-joint tree x 2000 frames scene (BASELINE.json configs[4]'s scene, every free parameter)",
            "free_parameters": int(idx.size), "samples": int(W.N_s), "ms_per_gradient": ms, "gradients_per_s": 1e3 / ms, "analytic_path_used": bool(used),
            "api": "omo_eval_grad_analytic (host x in, gradient out, wall clock)",
            "max_abs_diff_vs_central_differences_of_24_parameters": float(np.max(np.abs(g[sel] - g_fd))), "g_inf_of_those": float(np.max(np.abs(g_fd))),
            "finite_difference_columns_it_replaces": int(2 * idx.size + 1)}


def roofline_of(W, k_ms, n_se_executed, dfma_peak, step_ms):
    """F_se = 393 (bodies) + sum_bezier (45 + 14 k) [+ 92 J with positions] [+ 450 it_IK], with the curve
    and IK counters measured on the timed batch and only the WALKED bodies counted (candidate-independent subtrees are evaluated
    once per window, Appendix B: "report executed FLOPs")."""
    syn = _synthetic()
    eng = W.eng
    eng.set_counting(True)
    if W.grad:
        dX = eng.device_alloc(W.X.nbytes); dF = eng.device_alloc(8 * W.X.shape[0])
        eng.h2d(dX, W.X)
        eng.eval_device(dX, W.X.shape[0], dF, None, sync=True)
        eng.device_free(dX); eng.device_free(dF)
        n_counted = W.X.shape[0] * W.N_s
    else:
        eng.eval_device(W.dX_ptr, W.B_local, W.dF_ptr, None, sync=True)
        n_counted = W.B_local * W.N_s
    cnt = eng.counters()
    eng.set_counting(False)
    bez, newt, ikit = cnt["bezier_evals"] / n_counted, cnt["newton_iters"] / n_counted, cnt["ik_iters"] / n_counted
    walked, walked_general = eng.walked_bodies()
    flops_ref = syn.flops_per_sample_eval(W.J, bez, newt, with_positions=W.fk, ik_iters_per_sample=ikit)
    flops_se = syn.flops_per_sample_eval(walked + 1, bez, newt, with_positions=W.fk, ik_iters_per_sample=ikit)
    achieved = n_se_executed * flops_se / (k_ms * 1e-3)
    peaks = measured_peaks()
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    B_rows = W.X.shape[0] if W.grad else W.B_local
    alg_bytes = 8.0 * (B_rows * W.D + B_rows)                    # 8 (B D + B)
    cap = ncu_capture(W.name)
    jit = eng.jit_info()
    inst = eng.jit_describe()
    frac_exec = None
    if cap.get("fp64_flop_per_launch") and cap.get("sample_evals_per_launch"):
        frac_exec = cap["fp64_flop_per_launch"] / cap["sample_evals_per_launch"] * n_se_executed / (k_ms * 1e-3) / dfma_peak
    return {
        "bound": "fp64", "achieved": achieved / 1e12, "peak": dfma_peak / 1e12, "unit": "TFLOP/s", "frac": achieved / dfma_peak,
        "frac_definition": "ALGORITHMIC FLOPs of SURVEY.md section 8(d) on the bodies actually walked / kernel time / measured DFMA peak",
        "frac_executed": frac_exec,
        "frac_executed_definition": "FP64 FLOPs the kernel really retires (ncu: 2 DFMA + DMUL + DADD thread instructions, profiles/traffic.json) / kernel time / measured DFMA peak",
        "traffic": cap.get("dram_bytes_per_launch"), "traffic_source": cap.get("source"),
        "kernel": "omo_jit_kernel (scene-specialised, NVRTC)" if jit["launches"] > 0 else "omo::objective_kernel (table-driven)",
        "jit": {"launches": jit["launches"], "instances": inst, "log": jit["log"][:200]},
        "kernel_ms": k_ms, "kernel_share_of_step": k_ms / step_ms,
        "flops_per_sample_eval": flops_se, "flops_per_sample_eval_reference_work": flops_ref,
        "frac_of_reference_work": n_se_executed * flops_ref / (k_ms * 1e-3) / dfma_peak,
        "bodies": {"scene": W.J - 1, "walked": walked, "walked_general": walked_general},
        "bezier_evals_per_sample_eval": bez, "newton_iters_per_sample_eval": newt, "ik_iters_per_sample_eval": ikit,
        "transcendentals_per_sample_eval_not_counted": {"sincos_pairs": 3 * walked_general, "note": "SURVEY.md section 8(d) excludes them from F_se"},
        "peak_source": "dependent-free DFMA microbenchmark (omo_measure_dfma_peak) in this run; MEASURED_PEAKS.json has no FP64 figure",
        "hbm": {"algorithmic_bytes": alg_bytes, "algorithmic_GBps": alg_bytes / (k_ms * 1e-3) / 1e9, "peak_GBps": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"},
    }


def prepare_batch(G, W, first=0, B_local=None):
    """Device-resident and pinned-host copies of the workload's candidates."""
    syn = _synthetic()
    torch = G.torch
    eng = W.eng
    if W.grad:
        W.x_base = syn.perturbed_candidates(W.x0, 1, first=0)[0]
        W.X = fd_columns(W.x_base, W.idx)
        W.B_local = W.X.shape[0]
    else:
        W.B_local = W.B if B_local is None else B_local
        W.X = syn.perturbed_candidates(W.x0, W.B_local, first=first)
    W.dX = torch.empty((W.B_local, W.D), dtype=torch.float64, device="cuda")
    W.dF = torch.empty(W.B_local, dtype=torch.float64, device="cuda")
    W.dX_ptr, W.dF_ptr = W.dX.data_ptr(), W.dF.data_ptr()
    eng.h2d(W.dX_ptr, W.X)
    W.Xp = eng.pinned_array((W.B_local, W.D)); W.Xp[:] = W.X
    W.Fp = eng.pinned_array((W.B_local,))


def measure_single(G, W, steps, warmup, dfma_peak, cpu_candidates, full=False, sustain_s=0.0):
    """One workload on one GPU: device-resident value, e2e through the C ABI with host buffers, roofline, CPU parity."""
    eng = W.eng
    prepare_batch(G, W)
    B = W.B_local

    def step_device():
        if W.grad:   # the finite-difference columns through the API: built on the device, sparse over tiles
            eng.eval_grad_columns(W.x_base, W.idx, H, 0, B, W.dF_ptr, sync=False)
        else:
            eng.eval_device(W.dX_ptr, B, W.dF_ptr, None, sync=False)

    grad_out = {}

    def step_e2e():
        if W.grad:   # host x in, (f0, gradient, f+-) out on the host
            f0, g, fpm = eng.eval_grad(W.x_base, W.idx, H)
            grad_out["f0"], grad_out["g"], grad_out["fpm"] = f0, g, fpm
            W.Fp[0] = f0; W.Fp[1::2] = fpm[:, 0]; W.Fp[2::2] = fpm[:, 1]
        else:
            eng.eval_host_ptr(W.Xp.ctypes.data, B, W.Fp.ctypes.data)     # H2D of X, kernels, D2H of f, all inside

    # the step is timed WITHOUT the engine's own kernel-timing events in the stream; the objective kernel's time comes from a second pass
    ms_dev, wall_dev, _, launches = G.timed(eng, step_device, steps, warmup)
    _, _, (k_ms, k_n), _ = G.timed(eng, step_device, steps, 1, kernel_timing=True)
    ms_e2e, wall_e2e, _, _ = G.timed(eng, step_e2e, steps, max(warmup, 3))
    f_dev = W.dF.cpu().numpy()
    assert np.array_equal(f_dev, np.asarray(W.Fp)), "device-resident and host-buffer paths disagree"

    executed_frac = 1.0
    if W.grad:   # SURVEY.md section 8(d) dense counting rule: report the sample-evals actually executed
        ex_items, dense_items = eng.grad_sparsity()
        executed_frac = ex_items / max(dense_items, 1)
    n_se = B * W.N_s * executed_frac
    k_ms_per = k_ms / max(k_n, 1)
    roof = roofline_of(W, k_ms_per, n_se, dfma_peak, ms_dev)

    out = {"workload": W.desc, "value": n_se / (ms_dev * 1e-3), "unit": UNIT, "ms_per_step": ms_dev, "kernel_ms": k_ms_per,
           "e2e_ms_per_step": ms_e2e, "rows": B, "samples_per_row": W.N_s, "executed_fraction_of_dense_sample_evals": executed_frac,
           "roofline_frac": roof["frac"], "roofline_frac_executed": roof["frac_executed"], "kernel": roof["kernel"], "gpu_launches_per_step": launches / steps}
    if W.grad:
        out["objective_plus_gradient_evals_per_s"] = 1.0 / (ms_dev * 1e-3)
        out["e2e_objective_plus_gradient_evals_per_s"] = 1.0 / (ms_e2e * 1e-3)

    # ---- parity of what was timed against the CPU reference ----------------------------------------------------------
    cpu = None
    if cpu_candidates > 0:
        rows = W.X if W.grad else W.X[:cpu_candidates]
        leg = cpu_leg(W.name, W.scene, rows, as_is=full)
        rel = float(np.max(np.abs(f_dev[:rows.shape[0]] - leg["f"]) / np.abs(leg["f"])))
        assert rel <= 1e-9, f"{W.name}: GPU deviates from the CPU reference: {rel}"
        out["max_rel_diff_vs_oracle"] = rel
        out["oracle"] = {"kind": leg["kind"], "rows_compared": int(rows.shape[0]), "cpu_value": rows.shape[0] * leg["n_s"] / leg["seconds"], "cpu_cores": leg["cores"]}
        if W.grad and not W.core.item_names()[-1].endswith("IK"):
            # next row N4: the analytic gradient of the same point, timed through the Core API (host x in, gradient out)
            t_an = []
            for rep in range(8):
                t0 = time.perf_counter()
                f_an, g_an, used = W.core.CalculateAnalyticGradient(W.lo, W.hi, W.x_base, W.x0)
                t_an.append(time.perf_counter() - t0)
            g_fd_gpu = grad_out["g"]
            out["analytic_gradient"] = {"used": bool(used), "us_per_call": float(np.median(t_an[2:]) * 1e6),
                                        "max_abs_diff_vs_fd_gradient": float(np.max(np.abs(g_an - g_fd_gpu))), "g_inf": float(np.max(np.abs(g_fd_gpu))),
                                        "api": "omo_eval_grad_analytic through Core::CalculateFullObjectiveAndAnalyticGradient (wall clock)"}
        if W.grad:
            F = W.idx.size
            g_ref = (leg["f"][1::2] - leg["f"][2::2]) / (2.0 * H)
            bound = 1e-9 * max(float(np.max(np.abs(g_ref))), 1.0) + 2e-14 * abs(float(leg["f"][0])) / (2.0 * H)
            gd = float(np.max(np.abs(grad_out["g"] - g_ref)))
            assert gd <= bound, f"{W.name}: gradient deviates: {gd} > {bound}"
            out["gradient"] = {"free_parameters": int(F), "max_abs_diff_vs_oracle": gd, "bound": bound, "g_inf": float(np.max(np.abs(g_ref)))}
        cpu = leg
    return out, roof, cpu, SimpleNamespace(ms_dev=ms_dev, ms_e2e=ms_e2e, wall_dev=wall_dev, wall_e2e=wall_e2e, launches=launches, n_se=n_se, f_dev=f_dev,
                                           step_device=step_device, k_ms=k_ms_per)


def measure_solve(G, dfma_peak):
    """BASELINE.json configs[1]: the fox-tail solve (both tail joints selected, F = 27, LN_BOBYQA, 2000 evaluations) with the
    GPU callback; every callback value replayed on the CPU reference. Then the gradient-based solve of the same problem."""
    from optimo_b200.core import Core
    scene = load_scene("fox-tail")
    c = Core(scene, device=G.local)
    _KEEP.append(c)
    c.select_items(["tail_root", "tail_center"])
    x_init = c.GetDescriptor()
    c.UpdateOriginalParameters(x_init)
    c.record_solve_trace(True)
    c.PerformOptimization(False)           # warm: lazy allocations, specialised-kernel build, graph capture
    c.SetDescriptor(x_init)
    t0 = time.perf_counter()
    f_end, evals = c.PerformOptimization(False)
    wall = time.perf_counter() - t0
    Xf, Fv = c.solve_trace()
    idx = c.free_indices()
    O, o, kind = make_oracle("foxtail_solve", scene)
    o.select_items(["tail_root", "tail_center"])
    o.update_original(x_init)
    rows = np.tile(x_init, (Xf.shape[0], 1))
    rows[:, idx] = Xf
    t1 = time.perf_counter()
    sec, f_cpu = o.bench_batch(rows, 1, O.hardware_concurrency(kind))
    rel = float(np.max(np.abs(Fv - f_cpu) / np.abs(f_cpu)))
    assert rel <= 1e-9, f"fox-tail solve: a callback value deviates from the CPU reference: {rel}"
    # the serial CPU callback, as the reference runs it (one evaluation after another on the as-is schedule)
    n_cpu = min(64, rows.shape[0])
    sec_serial, _ = o.bench_batch(rows[:n_cpu], 0, O.hardware_concurrency(kind))
    n_s = c.engine().num_samples
    out = {"workload": WORKLOADS["foxtail_solve"][0], "value": evals * n_s / wall, "unit": UNIT, "ms_per_step": wall * 1e3, "step": "one full solve",
           "evaluations": int(evals), "us_per_callback": wall / max(evals, 1) * 1e6, "f_initial": float(Fv[0]), "f_final": float(f_end),
           "max_rel_diff_vs_oracle": rel, "oracle": {"kind": "reference" if kind == "ref" else "port", "rows_compared": int(rows.shape[0])},
           "cpu_us_per_callback_as_is": sec_serial / n_cpu * 1e6, "roofline_frac": None,
           "roofline_note": "latency-bound: one candidate x 245 samples per launch (2 thread blocks); frac is not meaningful",
           "graph_replays": c.engine().graph_replays(), "callbacks_served_by_the_batched_start": c.last_solve_batched}
    # next row N1: the gradient-based solve over the finite-difference gradient
    c.SetDescriptor(x_init)
    c.PerformOptimizationWithGradient(False, "LD_SLSQP")
    c.SetDescriptor(x_init)
    t0 = time.perf_counter()
    fg, eg = c.PerformOptimizationWithGradient(False, "LD_SLSQP")
    out["gradient_solve"] = {"algorithm": "LD_SLSQP", "ms": (time.perf_counter() - t0) * 1e3, "callbacks": int(eg), "f_final": float(fg)}
    c.SetDescriptor(x_init)
    c.PerformOptimizationWithGradient(False, "LD_SLSQP", analytic=True)
    c.SetDescriptor(x_init)
    t0 = time.perf_counter()
    fa, ea = c.PerformOptimizationWithGradient(False, "LD_SLSQP", analytic=True)
    out["analytic_gradient_solve"] = {"algorithm": "LD_SLSQP", "ms": (time.perf_counter() - t0) * 1e3, "callbacks": int(ea), "f_final": float(fa)}
    return out


def sustained(G, W, run, seconds, dfma_peak, n_se, k_share):
    """>= `seconds` of back-to-back steps (no flush, no host round trip between them) with the clock record of that interval:
    the burst figure is a 20 ms region, this is what the kernel holds once the SM clock settles under load."""
    torch = G.torch
    eng = W.eng
    stream = G.stream_of(eng)
    n = max(10, int(seconds / (run.ms_dev * 1e-3)))
    sampler = ClockSampler(G.local)
    with torch.cuda.stream(stream):
        stream.synchronize()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eng.kernel_timing(reset=1)
        e0.record(stream)
        for _ in range(n):
            run.step_device()
        e1.record(stream)
        stream.synchronize()
        k_ms, k_n = eng.kernel_timing(reset=-1)
    sampler.stop()
    ms = e0.elapsed_time(e1) / n
    return {"seconds": e0.elapsed_time(e1) * 1e-3, "steps": n, "ms_per_step": ms, "value": n_se / (ms * 1e-3), "kernel_ms": k_ms / max(k_n, 1), "clocks": sampler.summary()}


def run_single_gpu(G, args):
    main_name = args.workload
    if main_name == "foxtail_solve":
        line = measure_solve(G, None)
        print(json.dumps(line), flush=True)
        return
    W = build(main_name, G.local)
    dfma_peak = W.eng.dfma_peak()
    sampler = ClockSampler(G.local)
    sampler.start()
    cores = os.cpu_count() or 1
    cpu_n = 0 if args.no_cpu_baseline else (max(cores, min(W.B, 64 * cores)) if main_name != "tree64" else max(cores, 16))
    out, roof, cpu, run = measure_single(G, W, args.steps, args.warmup, dfma_peak, cpu_n, full=True)

    ms_free = None
    if not W.grad:   # the same candidates as their free entries only (omo_eval_free: expansion over x0 on the device)
        eng = W.eng
        Xfp = eng.pinned_array((W.B_local, W.idx.size)); Xfp[:] = W.X[:, W.idx]
        Fp2 = eng.pinned_array((W.B_local,))
        ms_free, _, _, _ = G.timed(eng, lambda: eng.eval_free_host_ptr(W.x0, W.idx, Xfp.ctypes.data, W.B_local, Fp2.ctypes.data), args.steps, max(args.warmup, 3))
        assert np.array_equal(run.f_dev, np.asarray(Fp2)), "free-parameter and full-descriptor host paths disagree"
        # pageable host buffers (what a std::vector caller has): the copy is staged by the driver
        Xpage = W.X.copy(); Fpage = np.zeros(W.B_local)
        ms_page, _, _, _ = G.timed(eng, lambda: eng.eval_host_ptr(Xpage.ctypes.data, W.B_local, Fpage.ctypes.data), max(3, args.steps // 2), 3)
        assert np.array_equal(run.f_dev, Fpage)
    sampler.stop()
    clocks = sampler.summary()

    sus = None
    if args.sustain > 0:
        sus = sustained(G, W, run, args.sustain, dfma_peak, run.n_se, roof["kernel_share_of_step"])
        sus["roofline_frac"] = roof["frac"] * (roof["kernel_ms"] / sus["kernel_ms"])
        roof["sustained"] = sus

    cpu_baseline = None
    if cpu is not None:
        cpu_val = cpu["candidates"] * cpu["n_s"] / cpu["seconds"]
        cpu_baseline = {"value": cpu_val, "unit": UNIT, "cores": cpu["cores"], "kind": cpu["kind"],
                        "sample": f"first {cpu['candidates']} of {W.B_local} rows, candidate-parallel over {cpu['cores']} host threads, {cpu['seconds']:.2f} s; RBDL on its SimpleMath backend (Eigen is not installed)",
                        "max_rel_diff_vs_gpu": out.get("max_rel_diff_vs_oracle"), "as_is": cpu.get("as_is")}

    workloads = {main_name: out}
    if not args.no_workloads and main_name == "stick240":
        for name in ("grad", "grad_stick", "grad_conductor", "tree64"):
            t0 = time.perf_counter()
            Wk = build(name, G.local)
            o2, r2, _, _ = measure_single(G, Wk, max(5, args.steps // 2), max(3, args.warmup), dfma_peak, 0 if args.no_cpu_baseline else (16 if name == "tree64" else 1))
            o2["roofline"] = {k: r2[k] for k in ("frac", "frac_executed", "achieved", "peak", "kernel", "kernel_ms", "flops_per_sample_eval", "bodies",
                                                 "bezier_evals_per_sample_eval", "newton_iters_per_sample_eval", "ik_iters_per_sample_eval", "traffic")}
            o2["jit"] = r2["jit"]["instances"]
            o2["bench_wall_s"] = time.perf_counter() - t0
            workloads[name] = o2
            del Wk
        t0 = time.perf_counter()
        workloads["foxtail_solve"] = measure_solve(G, dfma_peak)
        workloads["foxtail_solve"]["bench_wall_s"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        try:
            workloads["tree64_analytic_gradient"] = measure_analytic_tree64(G)
        except Exception as e:   # noqa: BLE001  (an extra line: never the reason the bench line is missing)
            workloads["tree64_analytic_gradient"] = {"error": str(e)[:300]}
        workloads["tree64_analytic_gradient"]["bench_wall_s"] = time.perf_counter() - t0

    B, N_s, D, J = W.B_local, W.N_s, W.D, W.J
    n_total = run.n_se
    line = {
        "metric": METRIC, "value": n_total / (run.ms_dev * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": run.ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": W.desc, "candidates": B, "samples_per_candidate": N_s, "joints": J, "descriptor": D,
                   "l2": "flushed between timed iterations (256 MiB memset)", "gather": "none"},
        "clocks": clocks,
        "e2e": {"value": n_total / (run.ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": run.ms_e2e,
                "api": "omo_eval_grad (host x in; f0, gradient, f+- out)" if W.grad else "omo_eval (pinned host X in, f out; chunked copy overlapped with the kernels)",
                "h2d_bytes_per_step": int((8 * D + 4 * W.idx.size) if W.grad else B * D * 8), "d2h_bytes_per_step": int(B * 8)},
        "e2e_free_parameters": None if ms_free is None else {
            "value": n_total / (ms_free * 1e-3), "unit": UNIT, "ms_per_step": ms_free, "api": "omo_eval_free (candidates as their F free entries, expanded on the device)",
            "h2d_bytes_per_step": int(B * W.idx.size * 8 + D * 8 + W.idx.size * 4), "d2h_bytes_per_step": int(B * 8)},
        "e2e_pageable": None if ms_free is None else {"value": n_total / (ms_page * 1e-3), "unit": UNIT, "ms_per_step": ms_page,
                                                       "api": "omo_eval with pageable (malloc'ed) host buffers"},
        "gpu_launches": int(run.launches),
        "roofline": roof, "cpu_baseline": cpu_baseline,
        "workloads": workloads,
        "executed_fraction_of_dense_sample_evals": out["executed_fraction_of_dense_sample_evals"],
        "objective_evals_per_s": B / (run.ms_dev * 1e-3), "frame_evals_per_s": n_total / (run.ms_dev * 1e-3) / 5.0,
        "objective_plus_gradient_evals_per_s": (1.0 / (run.ms_dev * 1e-3)) if W.grad else (B / (2 * W.idx.size + 1)) / (run.ms_dev * 1e-3),
        "wall_s": {"device_loop": run.wall_dev, "e2e_loop": run.wall_e2e},
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# N > 1: strong scaling of the one batch, weak scaling, sharded gradient, gather latencies, single-process multi-GPU
# ---------------------------------------------------------------------------------------------------------------------
def run_multi_gpu(G, args):
    from optimo_b200.core import Core, MultiEngine, grad_from_columns, shard_range
    torch, dist = G.torch, G.dist
    world, rank = G.world, G.rank
    name = args.workload if args.workload in ("stick240", "tree64") else "stick240"
    W = build(name, G.local)
    eng = W.eng
    syn = _synthetic()
    dfma_peak = eng.dfma_peak()
    B = W.B
    X_all = syn.perturbed_candidates(W.x0, B, first=0)          # the ONE batch; every rank can build it, each uses its slice only
    b0, nb = shard_range(B, rank, world)
    eng.set_batch_hint(B)    # the kernel the whole batch gets, whatever a rank's share
    pad = (B + world - 1) // world
    dXs = torch.empty((max(nb, 1), W.D), dtype=torch.float64, device="cuda")
    dFs = torch.zeros(pad, dtype=torch.float64, device="cuda")
    dAll = torch.empty(pad * world, dtype=torch.float64, device="cuda")
    eng.h2d(dXs.data_ptr(), X_all[b0:b0 + nb])
    Xp = eng.pinned_array((max(nb, 1), W.D)); Xp[:nb] = X_all[b0:b0 + nb]
    Fall = torch.empty(pad * world, dtype=torch.float64).pin_memory()

    def strong_device():
        eng.eval_device(dXs.data_ptr(), nb, dFs.data_ptr(), None, sync=False)
        dist.all_gather_into_tensor(dAll, dFs)

    def strong_e2e():     # host shard in (pinned), gathered f on rank 0's host
        eng.eval_host_in(Xp.ctypes.data, nb, dFs.data_ptr(), None, sync=False)
        dist.all_gather_into_tensor(dAll, dFs)
        if rank == 0:
            Fall.copy_(dAll, non_blocking=True)

    sampler = ClockSampler(G.local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_nccl, wall_dev, _, launches = G.timed(eng, strong_device, args.steps, args.warmup)
    _, _, (k_ms, k_n), _ = G.timed(eng, strong_device, args.steps, 1, kernel_timing=True)   # kernel time from a pass of its own: no timing events inside the measured step
    ms_e2e, wall_e2e, _, _ = G.timed(eng, strong_e2e, args.steps, max(args.warmup, 3))
    f_strong = np.concatenate([dAll.cpu().numpy()[r * pad: r * pad + shard_range(B, r, world)[1]] for r in range(world)])

    # ---- the same step with the gather fused into the kernel that forms f (omo_peer_gather_*: NVLink / NVSwitch peer stores + flags, no
    # NCCL call). Used for the headline when every rank could map its peers and the gathered values equal the NCCL path's bit for bit.
    peer = {"available": False}
    pg = None
    ok = 1
    try:
        from optimo_b200.core import PeerGather
        pg = PeerGather(eng, rank, world, B)
    except Exception as e:   # noqa: BLE001
        peer["error"] = str(e)[:300]
        pg, ok = None, 0
    handles = [None] * world
    dist.all_gather_object(handles, pg.handle if pg is not None else None)   # every rank takes part, whatever happened above
    if ok and all(hd is not None for hd in handles):
        try:
            pg.connect(handles)
        except Exception as e:   # noqa: BLE001
            peer["error"] = str(e)[:300]
            ok = 0
    else:
        ok = 0
    t_ok = torch.tensor([ok], dtype=torch.int32, device="cuda")
    dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
    ms_dev, gather_used = ms_nccl, "nccl all_gather of f inside the timed step"
    if int(t_ok.item()) == 1:
        def strong_device_peer():
            pg.eval(dXs.data_ptr(), nb, b0, sync=False)
        ms_peer, wall_peer, _, launches_p = G.timed(eng, strong_device_peer, args.steps, args.warmup)
        _, _, (k_ms_p, k_n_p), _ = G.timed(eng, strong_device_peer, args.steps, 1, kernel_timing=True)
        f_peer = eng.d2h(pg.values(), (B,))
        same = bool(np.array_equal(f_peer, f_strong)) and pg.error() == 0
        t_same = torch.tensor([1 if same else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(t_same, op=dist.ReduceOp.MIN)
        same = int(t_same.item()) == 1
        peer = {"available": True, "ms_per_step": ms_peer, "nccl_ms_per_step": ms_nccl, "bit_identical_to_nccl_path_on_every_rank": same,
                "api": "omo_peer_gather_eval: finish_kernel's sums stored into every rank's gathered array through cudaIpc-mapped peer memory, flags, wait -- one launch"}
        if same and ms_peer < ms_nccl:
            ms_dev, wall_dev, k_ms, k_n, launches = ms_peer, wall_peer, k_ms_p, k_n_p, launches_p
            gather_used = "NVLink / NVSwitch peer stores from the kernel that forms f (omo_peer_gather_eval); nccl all_gather measured beside it (peer_gather.nccl_ms_per_step)"
    elif "error" not in peer:
        peer["error"] = "another rank could not map its peers"

    # ---- weak scaling: every rank its own 4096 candidates (the r01 figure) ------------------------------------------
    Ww = W
    dXw = torch.empty((B, W.D), dtype=torch.float64, device="cuda")
    dFw = torch.empty(B, dtype=torch.float64, device="cuda")
    dAllw = torch.empty(B * world, dtype=torch.float64, device="cuda")
    eng.h2d(dXw.data_ptr(), syn.perturbed_candidates(W.x0, B, first=rank * B))

    def weak_device():
        eng.eval_device(dXw.data_ptr(), B, dFw.data_ptr(), None, sync=False)
        dist.all_gather_into_tensor(dAllw, dFw)
    ms_weak, _, _, _ = G.timed(eng, weak_device, args.steps, args.warmup)
    if sampler:
        sampler.stop()

    # ---- the gradient's 2F+1 columns sharded over the ranks ------------------------------------------------------------
    idx = W.idx
    P = 2 * idx.size + 1
    x_base = X_all[0]
    c0, nc = shard_range(P, rank, world)
    padc = (P + world - 1) // world
    dCs = torch.zeros(padc, dtype=torch.float64, device="cuda")
    dCall = torch.empty(padc * world, dtype=torch.float64, device="cuda")
    Call = torch.empty(padc * world, dtype=torch.float64).pin_memory()
    grad_box = {}

    def grad_sharded():    # host x in; f0 and the gradient on rank 0's host
        if nc > 0:
            eng.eval_grad_columns(x_base, idx, H, c0, nc, dCs.data_ptr(), sync=False)
        dist.all_gather_into_tensor(dCall, dCs)
        if rank == 0:
            Call.copy_(dCall, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            cols = np.concatenate([Call.numpy()[r * padc: r * padc + shard_range(P, r, world)[1]] for r in range(world)])
            grad_box["f0"], grad_box["g"] = grad_from_columns(cols, H)
    ms_grad, _, _, _ = G.timed(eng, grad_sharded, max(20, args.steps), max(args.warmup, 5), flush=False)
    grad_same = None
    ms_grad1 = None
    if rank == 0:
        f0_1, g_1, _ = eng.eval_grad(x_base, idx, H)
        grad_same = bool(f0_1 == grad_box["f0"] and np.array_equal(g_1, grad_box["g"]))
        assert grad_same, "sharded gradient differs from the one-GPU gradient"
    # one-GPU gradient call on every rank at the same time would not be the one-GPU figure: rank 0 alone times it
    if G.cpu_group is not None:
        dist.barrier(group=G.cpu_group)
    if rank == 0:
        stream = G.stream_of(eng)
        with torch.cuda.stream(stream):
            for _ in range(5):
                eng.eval_grad(x_base, idx, H)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(20):
                eng.eval_grad(x_base, idx, H)
            e1.record(stream); stream.synchronize()
            ms_grad1 = e0.elapsed_time(e1) / 20

    # ---- gather latencies: NCCL all-gather against a device-to-pinned-host copy ----------------
    gather = {}
    for n_d in (78, 468, 4096):
        per = (n_d + world - 1) // world
        src = torch.zeros(per, dtype=torch.float64, device="cuda")
        dst = torch.empty(per * world, dtype=torch.float64, device="cuda")
        hst = torch.empty(per, dtype=torch.float64).pin_memory()
        stream = G.stream_of(eng)
        with torch.cuda.stream(stream):
            for _ in range(10):
                dist.all_gather_into_tensor(dst, src)
            stream.synchronize(); G.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(200):
                dist.all_gather_into_tensor(dst, src)
            e1.record(stream); stream.synchronize()
            t_ag = torch.tensor([e0.elapsed_time(e1) / 200 * 1e3], dtype=torch.float64, device="cuda")
            dist.all_reduce(t_ag, op=dist.ReduceOp.MAX)
            t0 = time.perf_counter()
            for _ in range(200):
                hst.copy_(src, non_blocking=True); stream.synchronize()
            t_d2h = (time.perf_counter() - t0) / 200 * 1e6
        gather[f"{n_d}_doubles"] = {"nccl_all_gather_us": float(t_ag.item()), "d2h_to_pinned_and_sync_us": t_d2h}

    # ---- one host thread, N GPUs: the product's multi-device entry point (rank 0 alone; the other ranks wait on the host) ------
    single = None
    if G.cpu_group is not None:
        dist.barrier(group=G.cpu_group)
    if rank == 0 and not args.no_single_process:
        try:
            cm = Core(W.scene, devices=list(range(world)))
            _KEEP.append(cm)
            Fm = eng.pinned_array((B,)); Xm = eng.pinned_array((B, W.D)); Xm[:] = X_all
            m = None
            res = {}
            f_m = cm.CalculateFullObjectiveBatch(W.lo, W.hi, X_all[:64], W.x0)      # creates the engines, builds / loads the kernels
            m = cm.multi_engine()
            for label, mode in (("host", MultiEngine.GATHER_HOST), ("nccl", MultiEngine.GATHER_NCCL)):
                try:
                    m.set_gather(mode)
                except Exception as e:   # noqa: BLE001
                    res[label] = {"unavailable": str(e)[:200]}
                    continue
                for _ in range(3):
                    m.eval_host_ptr(Xm.ctypes.data, B, Fm.ctypes.data)
                t0 = time.perf_counter()
                for _ in range(args.steps):
                    m.eval_host_ptr(Xm.ctypes.data, B, Fm.ctypes.data)
                t_b = (time.perf_counter() - t0) / args.steps
                assert np.array_equal(np.asarray(Fm), f_strong), "multi-device batch differs from the sharded ranks"
                for _ in range(5):
                    gm = m.eval_grad(x_base, idx, H)
                t0 = time.perf_counter()
                for _ in range(50):
                    gm = m.eval_grad(x_base, idx, H)
                t_g = (time.perf_counter() - t0) / 50
                assert gm[0] == grad_box["f0"] and np.array_equal(gm[1], grad_box["g"]), "multi-device gradient differs"
                res[label] = {"batch_ms": t_b * 1e3, "batch_value": B * W.N_s / t_b, "gradient_us": t_g * 1e6, "bit_identical": True}
            del m
            single = {"api": "omo_create_multi / omo_multi_eval / omo_multi_eval_grad (one host thread, one worker per device; host wall clock, pinned X in, f out)",
                      "devices": world, "gather": res}
        except Exception as e:   # noqa: BLE001
            single = {"error": str(e)[:300]}
    if G.cpu_group is not None:
        dist.barrier(group=G.cpu_group)

    if rank == 0:
        k_ms_per = k_ms / max(k_n, 1)
        n_total = B * W.N_s
        W.B_local = nb; W.dX_ptr = dXs.data_ptr(); W.dF_ptr = dFs.data_ptr()
        roof = roofline_of(W, k_ms_per, nb * W.N_s, dfma_peak, ms_dev)
        line = {
            "metric": METRIC, "value": n_total / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": W.desc, "candidates_total": B, "candidates_per_gpu": nb, "samples_per_candidate": W.N_s, "joints": W.J, "descriptor": W.D,
                       "split": "omo_shard_range over candidates; every rank evaluates its slice", "gather": gather_used,
                       "l2": "flushed between timed iterations (256 MiB memset)"},
            "clocks": sampler.summary() if sampler else None,
            "e2e": {"value": n_total / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "api": "omo_eval_host_in (pinned host slice in) + nccl all_gather + one copy of all f to rank 0's pinned host",
                    "h2d_bytes_per_step": int(nb * W.D * 8), "d2h_bytes_per_step": int(pad * world * 8)},
            "gpu_launches": int(launches),
            "roofline": roof, "cpu_baseline": None,
            "weak": {"value": B * world * W.N_s / (ms_weak * 1e-3), "unit": UNIT, "ms_per_step": ms_weak, "candidates_per_gpu": B, "scaling": "weak"},
            "grad_sharded": {"columns": P, "columns_per_gpu": int(padc), "us_per_call": ms_grad * 1e3, "one_gpu_us_per_call": None if ms_grad1 is None else ms_grad1 * 1e3,
                             "api": "omo_eval_grad_columns(col_begin, col_count) per rank + nccl all_gather + omo_grad_from_columns on rank 0's host",
                             "bit_identical_to_one_gpu": grad_same},
            "gather_latency": gather,
            "peer_gather": peer,
            "single_process": single,
            "objective_evals_per_s": B / (ms_dev * 1e-3), "frame_evals_per_s": n_total / (ms_dev * 1e-3) / 5.0,
            "wall_s": {"device_loop": wall_dev, "e2e_loop": wall_e2e},
        }
        print(json.dumps(line), flush=True)
    if pg is not None:
        dist.barrier()      # nobody unmaps a peer that may still be written to
        pg.close()
    dist.destroy_process_group()


def run_ours(args):
    G = Gpu()
    if G.world > 1:
        run_multi_gpu(G, args)
    else:
        run_single_gpu(G, args)


def _reserve_stdout():
    """The contract is ONE JSON line on stdout. Libraries write there too (NCCL prints its version line on fd 1 when the first
    communicator comes up), so fd 1 is pointed at stderr for the run and the line goes to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(saved, "w")


def main():
    global print
    import faulthandler
    faulthandler.enable()
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
    ap.add_argument("--workload", default="stick240", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true", help="development runs: skip the timed CPU legs (cpu_baseline = null, no parity figures)")
    ap.add_argument("--no-workloads", action="store_true", help="only the main workload (profiling runs)")
    ap.add_argument("--no-single-process", action="store_true", help="N > 1: skip the omo_create_multi measurement")
    ap.add_argument("--sustain", type=float, default=3.0, help="seconds of back-to-back steps for the sustained figure (0 = skip)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

"""Kernel-time experiment harness: one workload (stick240 | tree64), device-resident batch, objective-kernel time from the
engine's own CUDA events (omo_kernel_timing), L2 flushed between iterations, plus a digest of f so that variants selected
through environment variables (OMO_JIT_SUB, OMO_JIT_FORM, OMO_JIT_MAXBLOCKS, ...) can be compared bit for bit.

  python profiles/exp_kernel.py tree64 [steps]  ->  one JSON line
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if os.environ.get("OMO_EXP_PKG_ROOT"):   # another build of the package (e.g. an export of an older commit) under the same harness
    sys.path.insert(0, os.path.abspath(os.environ["OMO_EXP_PKG_ROOT"]))

import bench  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "stick240"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    B = int(sys.argv[3]) if len(sys.argv) > 3 else bench.WORKLOADS[name][1]
    import torch
    W = bench.build(name, 0)
    eng = W.eng
    X = bench._synthetic().perturbed_candidates(W.x0, B)
    dX, dF = eng.device_alloc(X.nbytes), eng.device_alloc(8 * B)
    eng.h2d(dX, X)
    flush = torch.empty(256 * 1024 * 1024 // 8, dtype=torch.float64, device="cuda")
    for _ in range(3):
        eng.eval_device(dX, B, dF, None, sync=True)
    times, steps_ms = [], []
    sampler = bench.ClockSampler(0)
    sampler.start()
    for _ in range(8):   # ~10 ms of load so that the sampler sees the clocks the kernel runs at
        eng.eval_device(dX, B, dF, None, sync=True)
    for _ in range(steps):
        flush.zero_()
        torch.cuda.synchronize()
        eng.kernel_timing(reset=1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = torch.cuda.ExternalStream(eng.stream())
        e0.record(st)
        eng.eval_device(dX, B, dF, None, sync=True)
        e1.record(st)
        torch.cuda.synchronize()
        ms, n = eng.kernel_timing(reset=-1)
        times.append(ms / max(n, 1))
        steps_ms.append(e0.elapsed_time(e1))
    t_end = __import__("time").time() + 0.35
    while __import__("time").time() < t_end:   # keep the load up for a few sampler periods
        eng.eval_device(dX, B, dF, None, sync=True)
    sampler.stop()
    f = eng.d2h(dF, (B,))
    counts = None
    if os.environ.get("OMO_EXP_COUNT"):   # executed curve work of the batch (counting instance of the kernel; 64 candidates)
        eng.set_counting(True)
        nb = min(B, 64)
        eng.eval_device(dX, nb, dF, None, sync=True)
        c0 = eng.counters()
        eng.eval_device(dX, nb, dF, None, sync=True)
        c1 = eng.counters()
        eng.set_counting(False)
        n_se = nb * eng.num_samples
        counts = {k: (c1[k] - c0[k]) / n_se for k in c1 if k != "launches"}
    info = eng.jit_describe()
    env = {k: v for k, v in os.environ.items() if k.startswith("OMO_")}
    print(json.dumps({"workload": name, "B": B, "kernel_ms": float(np.median(times)), "kernel_ms_min": float(np.min(times)),
                      "step_ms": float(np.median(steps_ms)), "f_sha": hashlib.sha256(f.tobytes()).hexdigest()[:16], "clocks": sampler.summary(), "f0": float(f[0]), "per_sample_eval": counts,
                      "env": env, "jit": [{k: i.get(k) for k in ("form", "registers", "smem_bytes", "from_disk_cache")} for i in info]}), flush=True)
    eng.device_free(dX); eng.device_free(dF)


if __name__ == "__main__":
    main()

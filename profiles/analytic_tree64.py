#!/usr/bin/env python
"""BASELINE.json configs[4] scene (64-joint tree x 2000 frames, 113 967 free parameters): the analytic gradient (omo_eval_grad_analytic)
of one x, timed, and checked against finite-difference columns of a slice of the parameters (the full finite-difference gradient
would be 227 935 objective columns of 189 945 doubles each -- 346 GB of perturbed descriptors)."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optimo_b200 import synthetic
from optimo_b200.core import Core, grad_from_columns

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
scene = synthetic.tree_scene(64, n_frames)
c = Core(scene)
x0 = c.GetDescriptor()
x = synthetic.perturbed_candidates(x0, 1, first=1)[0]
lo, hi = c.window
idx = c.free_indices()
t0 = time.perf_counter(); f, g, used = c.CalculateAnalyticGradient(lo, hi, x, x0); t_first = time.perf_counter() - t0
ts = []
for _ in range(3):
    t0 = time.perf_counter(); f, g, used = c.CalculateAnalyticGradient(lo, hi, x, x0); ts.append(time.perf_counter() - t0)
# finite differences for a slice of the free parameters (every 1000th), one by one through the batch API
eng = c.engine()
sel = np.arange(0, idx.size, max(1, idx.size // 100))[:100]
X = np.tile(x, (2 * sel.size, 1))
for k, s in enumerate(sel):
    X[2 * k, idx[s]] += 1e-4
    X[2 * k + 1, idx[s]] -= 1e-4
t0 = time.perf_counter(); fx = c.CalculateFullObjectiveBatch(lo, hi, X, x0); t_fd = time.perf_counter() - t0
g_fd = (fx[0::2] - fx[1::2]) / 2e-4
err = float(np.max(np.abs(g[sel] - g_fd)))
print(json.dumps({"scene": f"tree64 x {n_frames} frames", "free_parameters": int(idx.size), "samples": eng.num_samples, "analytic_used": bool(used),
                  "analytic_first_call_s": t_first, "analytic_ms_per_call": float(np.median(ts) * 1e3), "f": f,
                  "fd_check": {"parameters_checked": int(sel.size), "max_abs_diff": err, "g_inf_of_checked": float(np.max(np.abs(g_fd))),
                               "fd_seconds_for_these_columns_as_dense_batch": t_fd,
                               "fd_full_gradient_projected_s": t_fd * (2 * idx.size + 1) / (2 * sel.size)}}))

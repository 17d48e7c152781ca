#!/usr/bin/env python
"""What the analytic gradient is for: OptiMo's optimisation on a scene no finite-difference or derivative-free method can touch --
BASELINE.json configs[4]'s 64-joint tree x 2000 frames, 113 967 free parameters -- solved with NLopt's LD_LBFGS over
omo_eval_grad_analytic (reverse mode), with the reference's bounds, tolerances and 2000-evaluation cap (core_opt.cpp:99-152).
    python profiles/solve_tree64_lbfgs.py [frames]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optimo_b200 import synthetic
from optimo_b200.core import Core

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
sys.stdout.flush()
result_fd = os.dup(1)
os.dup2(2, 1)      # nlopt-util prints its summary on fd 1
c = Core(synthetic.tree_scene(64, n_frames))
x0 = c.GetDescriptor()
lo, hi = c.window
f_start = c.CalculateFullObjective(lo, hi, x0, x0)
c.CalculateAnalyticGradient(lo, hi, x0, x0)      # allocations outside the timed solve
t0 = time.perf_counter()
f_end, evals = c.PerformOptimizationWithGradient(False, "LD_LBFGS", analytic=True)
wall = time.perf_counter() - t0
os.write(result_fd, (json.dumps({"scene": f"64-joint tree x {n_frames} frames", "free_parameters": int(c.free_indices().size), "samples": c.engine().num_samples,
                                 "algorithm": "LD_LBFGS over the analytic gradient", "f_start": f_start, "f_end": f_end, "callbacks": evals, "solve_s": wall,
                                 "ms_per_callback": wall / max(evals, 1) * 1e3}) + "\n").encode())

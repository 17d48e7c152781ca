#!/usr/bin/env python
"""Per-call time of the analytic gradient (omo_eval_grad_analytic through Core::CalculateAnalyticGradient) on the shipped scenes, reverse
mode (default) against forward mode (OMO_ADJ_MODE=forward), next to the one-launch finite-difference gradient of the same x."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optimo_b200.core import Core

out = {}
for name in ("stick-figure", "fox-tail"):
    row = {}
    for mode in ("reverse", "forward"):
        os.environ["OMO_ADJ_MODE"] = mode
        c = Core(os.path.join(ROOT, "tests", "golden", "scenes", name + ".json"))
        x0 = c.GetDescriptor()
        rng = np.random.default_rng(0)
        x = x0.copy(); x[2::5] += 0.3 * (rng.random(x.size // 5) - 0.5)
        lo, hi = c.window
        for _ in range(20):
            f, g, used = c.CalculateAnalyticGradient(lo, hi, x, x0)
        t0 = time.perf_counter()
        for _ in range(300):
            f, g, used = c.CalculateAnalyticGradient(lo, hi, x, x0)
        row[mode + "_us"] = (time.perf_counter() - t0) / 300 * 1e6
        row[mode + "_g_inf"] = float(np.max(np.abs(g)))
        if mode == "reverse":
            for _ in range(20):
                c.CalculateGradient(lo, hi, x, x0)
            t0 = time.perf_counter()
            for _ in range(300):
                gfd = c.CalculateGradient(lo, hi, x, x0)
            row["finite_difference_one_launch_us"] = (time.perf_counter() - t0) / 300 * 1e6
            row["max_abs_diff_vs_fd"] = float(np.max(np.abs(g - gfd[1])))
    out[name] = row
print(json.dumps(out, indent=1))

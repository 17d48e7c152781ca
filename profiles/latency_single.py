#!/usr/bin/env python
"""Latency of ONE objective evaluation through the callback API (Core::CalculateFullObjective -> omo_eval, B = 1), the call
NLopt makes per iteration (BASELINE.json configs[1]): the one-launch self-preparing kernel (default), the CUDA-graph replay of the
five-node chain (OMO_SELF=0) and plain launches (OMO_SELF=0 OMO_GRAPH=0), and
the wall time of the reference's BOBYQA set-up on fox-tail with either. Run on the GPU box:  python profiles/latency_single.py"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def per_call(scene, graph, n=3000, one_launch=False):
    os.environ["OMO_GRAPH"] = "1" if graph else "0"
    os.environ["OMO_SELF"] = "1" if one_launch else "0"
    from optimo_b200.core import Core
    c = Core(os.path.join(ROOT, "tests", "golden", "scenes", scene + ".json"))
    x0 = c.GetDescriptor()
    lo, hi = c.window
    rng = np.random.default_rng(0)
    X = x0 + 1e-3 * rng.standard_normal((16, x0.size)) * (np.arange(x0.size) % 5 == 2)
    for i in range(50):
        c.CalculateFullObjective(lo, hi, X[i % 16], x0)
    t0 = time.perf_counter()
    for i in range(n):
        c.CalculateFullObjective(lo, hi, X[i % 16], x0)
    us = (time.perf_counter() - t0) / n * 1e6
    for i in range(20):
        c.CalculateGradient(lo, hi, X[i % 16], x0)       # first call compiles the scene-specialised kernel (NVRTC)
    t0 = time.perf_counter()
    for i in range(n // 10):
        c.CalculateGradient(lo, hi, X[i % 16], x0)
    us_g = (time.perf_counter() - t0) / (n // 10) * 1e6
    return us, c.engine().graph_replays(), us_g, c.engine().jit_info()["compile_ms"]


def solve(graph, algorithm=None, one_launch=False):
    os.environ["OMO_GRAPH"] = "1" if graph else "0"
    os.environ["OMO_SELF"] = "1" if one_launch else "0"
    from optimo_b200.core import Core
    c = Core(os.path.join(ROOT, "tests", "golden", "scenes", "fox-tail.json"))
    c.select_items(["tail_root", "tail_center"])
    x = c.GetDescriptor()
    c.CalculateFullObjective(*c.window, x, x)
    if algorithm:
        c.CalculateGradient(*c.window, x, x)     # NVRTC outside the timed solve
    t0 = time.perf_counter()
    f, evals = c.PerformOptimizationWithGradient(False, algorithm) if algorithm else c.PerformOptimization(False)
    return (time.perf_counter() - t0) * 1e3, evals, f


if __name__ == "__main__":
    sys.stdout.flush()
    result_fd = os.dup(1)      # nlopt-util prints its own summary on fd 1: keep that out of the JSON
    os.dup2(2, 1)
    out = {}
    for scene in ("fox-tail", "stick-figure", "conductor"):
        a = per_call(scene, False)
        b = per_call(scene, True)
        s1 = per_call(scene, True, one_launch=True)
        out[scene] = {"plain_us_per_call": round(a[0], 2), "graph_us_per_call": round(b[0], 2), "graph_replays": b[1],
                      "one_launch_us_per_call": round(s1[0], 2),
                      "objective_plus_gradient_us_per_call": round(b[2], 2), "one_launch_objective_plus_gradient_us_per_call": round(s1[2], 2),
                      "nvrtc_compile_ms": b[3]}
    try:
        ms0, ev0, f0 = solve(False)
        ms1, ev1, f1 = solve(True)
        ms2, ev2, f2 = solve(True, one_launch=True)
        out["fox-tail BOBYQA solve"] = {"plain_ms": round(ms0, 1), "graph_ms": round(ms1, 1), "one_launch_ms": round(ms2, 1), "evals": [ev0, ev1, ev2], "f": [f0, f1, f2]}
        for alg in ("LD_SLSQP", "LD_LBFGS"):
            ms, ev, f = solve(True, alg)
            msb, evb, fb = solve(True, alg, one_launch=True)
            out[f"fox-tail {alg} solve"] = {"ms": round(ms, 1), "one_launch_ms": round(msb, 1), "callbacks": [ev, evb], "f": [f, fb]}
    except Exception as e:   # NLopt missing on the box
        out["fox-tail BOBYQA solve"] = {"error": str(e)}
    os.write(result_fd, (json.dumps(out, indent=1) + "\n").encode())

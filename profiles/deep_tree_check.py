"""A large batch on the deepest trees the C ABI admits (128 joints as one chain / as a binary tree): which kernel serves it
(scene-specialised or the deep-stack instance of the table-driven one), how long NVRTC took, and that both agree."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optimo_b200 import synthetic  # noqa: E402
from optimo_b200.core import Core  # noqa: E402

out = {}
for shape in ("binary", "chain"):
    scene = synthetic.tree_scene(n_joints=128, n_frames=12, key_every=4)
    if shape == "chain":
        nodes = list(synthetic.iter_joints(scene["joints"]))
        for n in nodes:
            n["children"] = []
        for a, b in zip(nodes, nodes[1:]):
            a["children"] = [b]
    res = {}
    for mode in ("0", "2"):
        os.environ["OMO_JIT"] = mode
        c = Core(scene)
        c.UpdateGlobalWeight(3.0)
        x0 = c.GetDescriptor()
        X = synthetic.perturbed_candidates(x0, 1024)
        lo, hi = c.window
        t0 = time.perf_counter()
        f = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
        first = time.perf_counter() - t0
        t0 = time.perf_counter()
        f = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
        second = time.perf_counter() - t0
        info = c.engine().jit_info()
        res[mode] = f
        out[f"{shape} OMO_JIT={mode}"] = {"first_call_s": round(first, 3), "second_call_ms": round(second * 1e3, 3), "jit_launches": info["launches"],
                                          "compile_ms": info["compile_ms"], "log": info["log"][:200]}
    out[f"{shape} max_rel_diff"] = float(np.max(np.abs(res["0"] - res["2"]) / np.abs(res["0"])))
print(json.dumps(out, indent=1))

import os, sys, time, json
import numpy as np
sys.path.insert(0, '/root/repo')
from optimo_b200.core import Core
out = {}
for scene in ("fox-tail", "stick-figure", "conductor"):
    c = Core(os.path.join('/root/repo', "tests", "golden", "scenes", scene + ".json"))
    if scene == "conductor":
        c.UpdateGlobalWeight(12.5)
    x0 = c.GetDescriptor(); lo, hi = c.window
    rng = np.random.default_rng(0)
    X = x0 + 1e-3 * rng.standard_normal((16, x0.size)) * (np.arange(x0.size) % 5 == 2)
    for i in range(50): c.CalculateFullObjective(lo, hi, X[i % 16], x0)
    n = 3000
    t0 = time.perf_counter()
    for i in range(n): c.CalculateFullObjective(lo, hi, X[i % 16], x0)
    us = (time.perf_counter() - t0) / n * 1e6
    eng = c.engine()
    # raw engine call (no Core sync per call)
    t0 = time.perf_counter()
    for i in range(n): eng.eval(X[i % 16])
    us_raw = (time.perf_counter() - t0) / n * 1e6
    eng.kernel_timing(reset=1)
    for i in range(200): eng.eval(X[i % 16])
    kms, kn = eng.kernel_timing(reset=-1)
    for i in range(20): c.CalculateGradient(lo, hi, X[i % 16], x0)
    t0 = time.perf_counter()
    for i in range(300): c.CalculateGradient(lo, hi, X[i % 16], x0)
    us_g = (time.perf_counter() - t0) / 300 * 1e6
    eng.kernel_timing(reset=1)
    for i in range(100): c.CalculateGradient(lo, hi, X[i % 16], x0)
    gms, gn = eng.kernel_timing(reset=-1)
    out[scene] = {"core_us": round(us, 1), "engine_us": round(us_raw, 1), "kernel_us": round(kms / max(kn, 1) * 1e3, 1), "grad_us": round(us_g, 1), "grad_kernel_us": round(gms / max(gn, 1) * 1e3, 1), "jit": [ (k['form'], k['self_preparing'], k['registers'], k['smem_bytes']) for k in eng.jit_describe()]}
print(json.dumps(out))

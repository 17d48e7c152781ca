import time, numpy as np, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from optimo_b200.core import Core
from conftest import scene_path
for name in ("fox-tail", "stick-figure"):
    c = Core(scene_path(name))
    if name == "fox-tail": c.select_items(["tail_root", "tail_center"])
    x0 = c.GetDescriptor(); lo, hi = c.window
    idx = c.free_indices()
    eng = c.engine(synced_window=(lo, hi))
    for fn, label in ((lambda: eng.eval_grad(x0, idx, 1e-4), "eng.eval_grad"), (lambda: c.CalculateGradient(lo, hi, x0, x0), "Core.CalculateGradient"),
                      (lambda: c.CalculateFullObjective(lo, hi, x0, x0), "Core.CalculateFullObjective")):
        for _ in range(5): fn()
        t = time.perf_counter()
        for _ in range(200): fn()
        print(name, label, (time.perf_counter() - t) / 200 * 1e6, "us", eng.grad_sparsity(), idx.size)

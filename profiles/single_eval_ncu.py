#!/usr/bin/env python
"""One scene, 60 single-candidate evaluations through the one-launch path -- the target of an `ncu --set full` capture of the
self-preparing kernel (run under ncu with --launch-skip 50 -c 2 -k regex:omo_jit_kernel)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optimo_b200.core import Core
scene = sys.argv[1] if len(sys.argv) > 1 else "stick-figure"
c = Core(os.path.join(ROOT, "tests", "golden", "scenes", scene + ".json"))
x0 = c.GetDescriptor(); lo, hi = c.window
eng = c.engine(synced_window=(lo, hi))
for i in range(60):
    eng.eval(x0 + 1e-4 * i)
print("ok")

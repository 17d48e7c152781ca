"""Four single-candidate evaluations per shipped scene -- run under `ncu --metrics gpu__time_duration.sum` for the launch list of one callback."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from optimo_b200.core import Core
for s in ("fox-tail", "stick-figure", "conductor"):
    c = Core(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "scenes", s + ".json"))
    x0 = c.GetDescriptor(); lo, hi = c.window
    for i in range(4):
        c.CalculateFullObjective(lo, hi, x0, x0)

#!/usr/bin/env python
"""One-launch single-candidate calls (Core::CalculateFullObjective -> omo_eval, B = 1) and small gradients with the host waiting on
the output slots in mapped host memory (default: sentinel polling) against waiting on the stream (OMO_SPIN=0); the fox-tail BOBYQA solve with
either. Run on the GPU box:  python profiles/latency_spin.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "profiles"))
import latency_single as L  # noqa: E402

if __name__ == "__main__":
    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)
    out = {}
    for spin in ("1", "0", "1", "0"):
        os.environ["OMO_SPIN"] = spin
        for scene in ("fox-tail", "stick-figure", "conductor"):
            us, _, us_g, _ = L.per_call(scene, True, n=4000, one_launch=True)
            out.setdefault(scene, {}).setdefault("polling the outputs in mapped memory" if spin == "1" else "stream synchronize", []).append(
                {"objective_us": round(us, 2), "objective_plus_gradient_us": round(us_g, 2)})
        ms, ev, f = L.solve(True, one_launch=True)
        out.setdefault("fox-tail BOBYQA solve", {}).setdefault("polling the outputs in mapped memory" if spin == "1" else "stream synchronize", []).append(
            {"ms": round(ms, 1), "evals": ev, "f": f})
    os.write(result_fd, (json.dumps(out, indent=1) + "\n").encode())

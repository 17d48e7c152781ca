"""Golden fixture of the next rows N2 / N3: tests/golden/states_at.npz (authoring container only).

    python tests/golden/make_golden_states.py

Per shipped scene, from the reference-linked oracle (oracle/_ref: the reference's own RBDL v2.5.0 / curve.cpp objects
under the restated glue): Core::CalculateTorques(t) with SetCurrentState's q / qd / qdd (core_rbdl.cpp:168-198,264-289)
and Joint::GetEulerAngle(t) of every joint (what Core::ExportMotion writes, core_io.cpp:252-294), at every frame of
the scene plus off-grid, out-of-range, unsorted and closer-than-2dt times.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import oracle as O  # noqa: E402
from make_golden import SCENES, perturb  # noqa: E402


def times_for(scene):
    frames = np.arange(scene["min_frame"], scene["max_frame"] + 1, dtype=np.float64)
    extra = np.array([7.37, 7.47, 7.40, -3.0, scene["max_frame"] + 7.25, 12.000000001, 19.9, 20.1, 20.0, 20.0])
    return np.concatenate([frames, extra])


def main():
    assert O.available("ref"), "build oracle/_ref first: make -C oracle ref"
    out = {}
    for name in SCENES:
        with open(os.path.join(HERE, "scenes", f"{name}.json")) as fh:
            scene = json.load(fh)
        o = O.Oracle(scene, kind="ref")
        x0, x = perturb(o, seed=1)
        t = times_for(scene)
        for tag, xx in (("orig", x0), ("pert", x)):
            s = o.torques_at(xx, t)
            out[f"{name}/{tag}/x"] = xx
            out[f"{name}/{tag}/times"] = t
            for k, v in s.items():
                out[f"{name}/{tag}/{k}"] = v
            out[f"{name}/{tag}/euler"] = o.euler_angles_at(xx, t)
            print(name, tag, "M =", t.size, "|tau|_max =", np.abs(s["tau"]).max())
    np.savez_compressed(os.path.join(HERE, "states_at.npz"), **out)


if __name__ == "__main__":
    main()

"""Generates the committed golden fixtures of tests/golden/ (run in the authoring container only).

    python tests/golden/make_golden.py

Needs /root/reference (scene JSON inputs) and the reference-linked oracle oracle/_ref/liboracle_ref.so
(`make -C oracle ref`), i.e. the reference's own RBDL v2.5.0 / curve.cpp / parallel-util objects under the
restated OptiMo glue. The GPU box has neither, which is why the outputs are committed:

    scenes/<name>.json        the three shipped scenes (inputs; re-serialised compactly, values untouched)
    objective_<name>.npz      per-scene cases: x, x_original, settings -> f, (C, D_param, D_traj), per-sample
                              q / qd / qdd / tau / local_cost, positions, FD gradient (f+, f-, g)
    rbdl_known_answers.json   RBDL / curve known answers recorded from real reference objects
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import oracle as O  # noqa: E402

REF_RES = "/root/reference/resources"
SCENES = ["stick-figure", "fox-tail", "conductor"]

# Settings of SURVEY.md §8(d) configs 1-3
CASES = {
    "stick-figure": [
        dict(tag="wt0", w_traj=0.0, kernels=[], select=None, cyclic=False, per_item={}),
        dict(tag="wt12", w_traj=12.5, kernels=[], select=None, cyclic=False, per_item={}),
        dict(tag="cyc", w_traj=12.5, kernels=[(1.0, 20.0, 2.6)], select=None, cyclic=True, per_item={"left_arm": 3.0, "spine": 0.5}),
    ],
    "fox-tail": [
        dict(tag="sel", w_traj=12.5, kernels=[], select=["tail_root", "tail_center"], cyclic=False, per_item={}),
    ],
    "conductor": [
        dict(tag="ik", w_traj=12.5, kernels=[(1.0, 20.0, 2.6), (1.0, 10.0, 1.0)], select=None, cyclic=False, per_item={}),
        dict(tag="ik0", w_traj=0.0, kernels=[], select=None, cyclic=False, per_item={}),
    ],
}


def perturb(o, seed):
    """x = x_original + delta on the free parameters (theta, t_b, t_f), keeping every time-Bezier monotone."""
    rng = np.random.default_rng(seed)
    x0 = o.descriptor()
    x = x0.copy()
    n = x.size // 5
    x[2::5] += 0.3 * (rng.random(n) - 0.5)
    x[3::5] = -(0.5 + rng.random(n))
    x[4::5] = +(0.5 + rng.random(n))
    return x0, x


def main():
    assert O.available("ref"), "build oracle/_ref first: make -C oracle ref"
    os.makedirs(os.path.join(HERE, "scenes"), exist_ok=True)
    for name in SCENES:
        with open(os.path.join(REF_RES, f"scene-{name}.json")) as fh:
            scene = json.load(fh)
        with open(os.path.join(HERE, "scenes", f"{name}.json"), "w") as fh:
            json.dump(scene, fh, separators=(",", ":"))
        out = {}
        for case in CASES[name]:
            o = O.Oracle(scene, kind="ref")
            o.set_sampling_concurrency(1)
            x0, x = perturb(o, seed=1)
            o.set_weight_trajectory(case["w_traj"])
            o.set_kernels(case["kernels"])
            o.set_cyclic(case["cyclic"])
            for item, w in case["per_item"].items():
                o.set_per_item_weight(item, w)
            if case["select"] is not None:
                o.select_items(case["select"])
            o.update_original(x0)
            f, terms = o.full_objective(x, x0)
            cache = o.cache()
            p = o.positions(x)
            o.set_cyclic(case["cyclic"])
            f0, g, fpm = o.gradient(x, h=1e-4)
            t = case["tag"]
            out[f"{t}/x"] = x
            out[f"{t}/x_original"] = x0
            out[f"{t}/f"] = np.array(f)
            out[f"{t}/terms"] = terms
            out[f"{t}/free_idx"] = o.free_indices()
            out[f"{t}/grad"] = g
            out[f"{t}/fpm"] = fpm
            out[f"{t}/f0"] = np.array(f0)
            out[f"{t}/positions"] = p
            out[f"{t}/wparam"] = o.parameter_weights()
            for k, v in cache.items():
                out[f"{t}/{k}"] = v
            print(name, t, "f =", repr(f), "terms =", terms, "F =", g.size, "N_s =", cache["time"].size)
        with open(os.path.join(HERE, f"cases_{name}.json"), "w") as fh:
            json.dump(CASES[name], fh, indent=1)
        np.savez_compressed(os.path.join(HERE, f"objective_{name}.npz"), **out)

    # Known answers recorded from real RBDL v2.5.0 / curve.cpp objects,
    # re-derived here from the reference-linked build so that the file is reproducible.
    I = np.diag([0.01, 0.002, 0.01])
    q = 0.1 * np.arange(9)
    qd = 0.05 * np.arange(9) - 0.1
    qdd = 0.3 - 0.02 * np.arange(9)
    tau = O.probe_inverse_dynamics(
        "ref", parent=[0, 1, 2], jkind=[0, 1, 1], xt=[[0, -0.6, 0], [0, -0.6, 0], [0, 0.2, 0]], mass=[0.0, 2.0, 5.0],
        com=[[0, 0, 0], [0, 0.1, 0], [0, 0.25, 0]], inertia=[np.eye(3), I, I], gravity=[0, 0, 0], q=q, qd=qd, qdd=qdd)
    ok, qres, iters = O.probe_inverse_kinematics(
        "ref", xt=[[0.5, 0, 0], [0.5, 0, 0]], mass=[1.0, 1.0], com=[[0.25, 0, 0], [0.25, 0, 0]], inertia=[I, I], point=[0.5, 0, 0],
        target=[0.9, 0.6, 0.2])
    cps = [[4, -np.pi / 4, 0, -1, 1], [12, 0.4 * np.pi, 0, -1, 1], [20, 0.0, 0, -1, 1], [26, 0.4 * np.pi, 0, -1, 1], [36, -np.pi / 4, 0, -1, 1]]
    ts = [3.0, 4.0, 7.4000000000000021, 19.998999999999999, 30.199999999999999, 36.0, 41.800000000000054]
    vals = O.probe_curve("ref", cps, ts)
    known = {
        "inverse_dynamics_A2": {"tau": [float(v) for v in tau],
                                "survey": [2.7900015437343511, 1.0138351782258836, 1.623340752656889, -0.03649681155412296,
                                           0.78175963704318407, -0.56409663290348233, 0.15862270537220366, 0.41570770191453793,
                                           -0.23302090588456958]},
        "inverse_kinematics_A3": {"converged": bool(ok), "q": [float(v) for v in qres],
                                  "survey": [0.27528231208255421, -0.07000486108586966, -0.001423387056454417, 1.4411196811754616,
                                             -0.3300459055072173, -1.2298697794430612e-18]},
        "curve_A5": {"cps": cps, "t": ts, "value": [float(v) for v in vals],
                     "survey": [-0.78539816339744828, -0.78539816339744839, 0.060807641590248485, 4.1739623380780448e-07,
                                0.41694273320812641, -0.78539816339744828, -0.78539816339744828]},
    }
    for k, v in known.items():
        got = np.array(v["tau"] if "tau" in v else v["q"] if "q" in v else v["value"])
        assert np.allclose(got, v["survey"], rtol=1e-13, atol=1e-15), (k, got, v["survey"])
    with open(os.path.join(HERE, "rbdl_known_answers.json"), "w") as fh:
        json.dump(known, fh, indent=1)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()

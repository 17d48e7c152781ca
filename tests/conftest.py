"""Shared fixtures. `-m "not gpu"` runs here (no GPU); `-m gpu` runs on a B200 through the C ABI."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SCENES = ["stick-figure", "fox-tail", "conductor"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def scene_path(name):
    return os.path.join(GOLDEN, "scenes", f"{name}.json")


def load_scene(name):
    with open(scene_path(name)) as fh:
        return json.load(fh)


def load_cases(name):
    with open(os.path.join(GOLDEN, f"cases_{name}.json")) as fh:
        cases = json.load(fh)
    data = np.load(os.path.join(GOLDEN, f"objective_{name}.npz"))
    return cases, data


def all_cases():
    out = []
    for name in SCENES:
        cases, _ = load_cases(name)
        out += [(name, c["tag"]) for c in cases]
    return out


@pytest.fixture(scope="session")
def oracle_port():
    """The CPU oracle (restatement build); compiled on demand -- test infrastructure only."""
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
    from oracle import oracle as O
    return O


def configure(obj, case, is_oracle):
    """Applies one golden case's settings to an Oracle or a product Core."""
    if is_oracle:
        obj.set_weight_trajectory(case["w_traj"])
        obj.set_kernels([tuple(k) for k in case["kernels"]])
        obj.set_cyclic(case["cyclic"])
        for item, w in case["per_item"].items():
            obj.set_per_item_weight(item, w)
        if case["select"] is not None:
            obj.select_items(case["select"])
    else:
        obj.UpdateGlobalWeight(case["w_traj"])
        obj.set_kernels([tuple(k) for k in case["kernels"]])
        obj.set_cyclic(case["cyclic"])
        for item, w in case["per_item"].items():
            obj.UpdatePerItemWeight(w, item)
        if case["select"] is not None:
            obj.select_items(case["select"])


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


def elem_err(a, b, rtol):
    """Element-wise deviation in units of the allowed one: max_i |a_i - b_i| / (rtol * |b_i| + 1e-12 * max|b|).
    <= 1 passes. Unlike rel_err (one ratio for the whole array) a small entry cannot hide behind a large one: an entry of
    1e-6 of the largest that is 100 % wrong scores ~1e6. The absolute floor (1e-12 of the largest entry) is what SURVEY.md
    App. B prescribes for the O(1e-16) stencil residues of constant channels, which the pruned walk returns as exact zeros."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    floor = 1e-12 * max(float(np.max(np.abs(b))) if b.size else 0.0, 1e-300)
    return float(np.max(np.abs(a - b) / (rtol * np.abs(b) + floor))) if b.size else 0.0


def gradient_bound(g_ref, f0, h, tol=1e-9):
    """Allowed |g - g_ref|_inf. SURVEY.md section 8(d): 1e-9 * max(|g|_inf, 1). The central difference divides two values that
    each carry the summation-order difference between the tile-tree reduction and the reference's serial loop (<= 2e-14
    relative, N_s terms) by 2h, so that share is added explicitly (DESIGN.md section 2 states the same bound)."""
    return tol * max(float(np.max(np.abs(g_ref))) if g_ref.size else 0.0, 1.0) + 2e-14 * abs(f0) / (2.0 * h)

"""Entry points of include/omo.h that round 2 added around the hot path: host-in / device-out evaluation (what a rank of a
multi-process job calls before its NCCL gather), the batch hint (kernel choice by the WHOLE batch), the description of the
specialised instances, and the on-disk cubin cache."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, load_scene, scene_path

from optimo_b200 import synthetic
from optimo_b200.core import Core

pytestmark = pytest.mark.gpu


def test_host_in_device_out_equals_the_host_call():
    c = Core(synthetic.stick_240(load_scene("stick-figure")))
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    for B in (3, 700):                                  # 700 x 780 doubles = 4.4 MB: the chunked, two-stream path
        X = synthetic.perturbed_candidates(x0, B)
        f_host, t_host = eng.eval(X, want_terms=True)
        Xp = eng.pinned_array((B, x0.size)); Xp[:] = X
        dF, dT = eng.device_alloc(8 * B), eng.device_alloc(24 * B)
        eng.eval_host_in(Xp.ctypes.data, B, dF, dT, sync=True)
        assert np.array_equal(eng.d2h(dF, (B,)), f_host) and np.array_equal(eng.d2h(dT, (B, 3)), t_host)
        eng.device_free(dF); eng.device_free(dT)


def test_batch_hint_makes_a_slice_bit_identical_to_its_rows_in_the_whole_batch():
    """A device of omo_multi / a rank of a multi-process job evaluates a SLICE; with the hint it takes the kernel the whole batch
    would (scene-specialised from 64 K sample-evals up), so the split never changes a bit."""
    c = Core(synthetic.stick_240(load_scene("stick-figure")))
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    X = synthetic.perturbed_candidates(x0, 256)          # 256 x 1216 = 311 K sample-evals: specialised kernel
    f_all = eng.eval(X)
    assert eng.jit_info()["launches"] > 0
    small = eng.eval(X[10:14])                           # 4 rows on their own: the table-driven kernel
    assert np.allclose(small, f_all[10:14], rtol=1e-12, atol=0)
    eng.set_batch_hint(256)
    assert np.array_equal(eng.eval(X[10:14]), f_all[10:14])
    assert np.array_equal(eng.eval(X[200:201]), f_all[200:201])      # even one row stays on the batch kernels
    eng.set_batch_hint(0)
    inst = eng.jit_describe()
    assert inst and all(k in inst[0] for k in ("form", "registers", "smem_bytes", "build_ms", "from_disk_cache", "self_preparing"))


def test_on_disk_cubin_cache_serves_a_second_process(tmp_path):
    """A specialised kernel costs seconds of NVRTC time; the second process on the same scene loads the cubin from
    $OMO_JIT_CACHE_DIR instead (keyed by the generated source and the embedded headers), with identical results."""
    code = (
        "import json, sys; sys.path.insert(0, %r)\n"
        "from optimo_b200.core import Core\n"
        "c = Core(%r); x = c.GetDescriptor(); lo, hi = c.window\n"
        "f0, g, _ = c.CalculateGradient(lo, hi, x, x)\n"
        "print(json.dumps({'f0': f0, 'g0': float(g[0]), 'inst': c.engine().jit_describe()}))\n") % (ROOT, scene_path("fox-tail"))
    env = dict(os.environ, OMO_JIT_CACHE_DIR=str(tmp_path / "cache"))
    runs = [json.loads(subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout.strip().splitlines()[-1])
            for _ in range(2)]
    assert not any(i["from_disk_cache"] for i in runs[0]["inst"]) and all(i["from_disk_cache"] for i in runs[1]["inst"])
    assert runs[0]["f0"] == runs[1]["f0"] and runs[0]["g0"] == runs[1]["g0"]
    assert any(fn.endswith(".cubin") for fn in os.listdir(tmp_path / "cache"))

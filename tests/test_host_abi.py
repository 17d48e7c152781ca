"""CPU-side checks of the product: the C-ABI libraries load and export every symbol the headers declare,
the host mirror of the reference's Core/Object (JSON, descriptor, physical parameters, weights, frozen
dofs) agrees with the oracle, and the compute path fails LOUDLY without a GPU (no CPU fallback)."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

from conftest import ROOT, SCENES, load_scene, scene_path

import optimo_b200
from optimo_b200 import synthetic
from optimo_b200.core import Core, EngineError, grad_from_columns, shard_range


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:omo|optimo)_[a-z0-9_]+)\s*\(", src)))


@pytest.mark.parametrize("header,lib", [("omo.h", "omo"), ("optimo_core.h", "core")])
def test_library_exports_every_declared_symbol(header, lib):
    names = _declared(header)
    assert len(names) > 20
    handle = ctypes.CDLL(optimo_b200.lib_paths()[lib])
    missing = [n for n in names if not hasattr(handle, n)]
    assert not missing, missing


def test_no_oracle_dependency_in_product():
    """The product must never import, link or call oracle/ (checked on sources and on the linked libraries)."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "optimo_b200")):
        for fn in files:
            if fn.endswith((".py", ".cpp", ".hpp", ".cu", ".cuh", ".h", "Makefile")):
                text = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "liboracle" not in text and "oracle/" not in text and "import oracle" not in text, os.path.join(dirpath, fn)
    for p in optimo_b200.lib_paths().values():
        assert b"liboracle" not in open(p, "rb").read()


@pytest.mark.parametrize("name", SCENES)
def test_scene_json_flattening_matches_oracle(oracle_port, name):
    o = oracle_port.Oracle(load_scene(name), "port")
    c = Core(scene_path(name))
    assert c.J == o.J and c.D == o.D
    assert c.item_names() == o.item_names
    assert np.array_equal(c.GetDescriptor(), o.descriptor())
    assert np.array_equal(c.freezed_dofs_, o.freezed())
    for j in range(c.J):
        m, com, I = c.joint_physical(j)
        mo, como, Io = o.joint_physical(j)
        assert m == mo and np.array_equal(com, como)
        assert np.allclose(I, Io, rtol=1e-13, atol=1e-17)
    # descriptor round trip and cyclic tangent averaging (object.cpp:147-186)
    x = c.GetDescriptor() + 0.01
    c.SetDescriptor(x)
    assert np.array_equal(c.GetDescriptor(), x)
    c.set_cyclic(True); o.set_cyclic(True)
    c.SetDescriptor(x); o.lib.orc_set_descriptor(o.h, x.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), x.size)
    assert np.array_equal(c.GetDescriptor(), o.descriptor())


def test_weights_selection_and_kernels_match_oracle(oracle_port):
    o = oracle_port.Oracle(load_scene("conductor"), "port")
    c = Core(scene_path("conductor"))
    for obj, setter in ((o, lambda n, w: o.set_per_item_weight(n, w)), (c, lambda n, w: c.UpdatePerItemWeight(w, n))):
        setter("Spine", 2.5); setter("LeftArmIK", 0.75)
    o.set_weight_trajectory(12.5); c.UpdateGlobalWeight(12.5)
    assert np.array_equal(c.CalculateParameterWeights(), o.parameter_weights())
    o.select_items(["Spine", "LeftArmIK"]); c.select_items(["Spine", "LeftArmIK"])
    assert np.array_equal(c.freezed_dofs_, o.freezed())
    ks = [(1.0, 20.0, 2.6), (1.0, 10.0, 1.0)]
    o.set_kernels(ks); c.set_kernels(ks)
    for t in (-2.0, 9.3, 10.0, 19.0, 21.7, 42.0):
        assert c.GetCostWeight(t) == o.cost_weight(t)
    with pytest.raises(EngineError):
        c.UpdatePerItemWeight(1.0, "no-such-item")


def test_default_humanoid_is_the_stick_figure_scene():
    """Core::Core() (core.cpp:23-103) builds the humanoid that scene-stick-figure.json stores."""
    a, b = Core(), Core(scene_path("stick-figure"))
    assert a.item_names() == b.item_names()
    assert np.allclose(a.GetDescriptor(), b.GetDescriptor(), rtol=0, atol=1e-15)


def test_json_write_read_round_trip(tmp_path):
    c = Core(scene_path("conductor"))
    p = tmp_path / "scene.json"
    c.WriteJson(p)
    d = Core(p)
    assert d.item_names() == c.item_names() and np.array_equal(d.GetDescriptor(), c.GetDescriptor())
    assert json.loads(c.scene_json())["ik_handlers"][0]["root_name"] == "LeftArm"
    with pytest.raises(EngineError):
        Core().ReadJsonString("{ not json")


def test_shard_range_and_gradient_assembly():
    for n, world in ((937, 8), (157, 4), (5, 8), (0, 2), (64, 1)):
        spans = [shard_range(n, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == n
        for (b0, c0), (b1, _) in zip(spans, spans[1:]):
            assert b1 == b0 + c0
        assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    cols = np.array([5.0, 7.0, 3.0, 2.0, 2.5])
    f0, g = grad_from_columns(cols, 0.25)
    assert f0 == 5.0 and np.array_equal(g, [(7.0 - 3.0) / 0.5, (2.0 - 2.5) / 0.5])


def test_synthetic_generators():
    rng = synthetic.MT19937_64([5489])
    assert int(rng.next_u64()[0]) == 14514284786278117030   # std::mt19937_64 default-seed first output
    assert int(rng.next_u64()[0]) == 4620546740167642908
    stick = load_scene("stick-figure")
    big = synthetic.stick_240(stick)
    assert synthetic.count_control_points(big) == 156 and big["max_frame"] == 240
    c = Core(big)
    assert c.D == 780 and c.free_indices().size == 468
    X = synthetic.perturbed_candidates(c.GetDescriptor(), 5)
    assert X.shape == (5, 780) and np.array_equal(X[:, 0::5], np.tile(c.GetDescriptor()[0::5], (5, 1)))
    assert (X[:, 3::5] <= -0.5).all() and (X[:, 4::5] >= 0.5).all()
    assert np.array_equal(X[3], synthetic.perturbed_candidates(c.GetDescriptor(), 1, first=3)[0])
    tree = synthetic.tree_scene()
    assert synthetic.count_control_points(tree) == 37989
    assert Core(tree).J == 64


def test_compute_fails_loudly_without_a_gpu():
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present; the loud-failure path is for CPU-only hosts")
    c = Core(scene_path("stick-figure"))
    x = c.GetDescriptor()
    with pytest.raises(EngineError, match="no usable CUDA device"):
        c.CalculateFullObjective(-1.0, 42.0, x, x)


def test_history_stack_and_files_need_no_gpu(tmp_path):
    """Next row N3: PushHistory / PopHistory / WriteHistory (core_io.cpp:171-213) are host-side text handling."""
    c = Core(scene_path("conductor"))
    x = c.GetDescriptor()
    c.PushHistory()
    s0 = c.scene_json()
    x2 = x.copy()
    x2[2::5] -= 0.05
    c.SetDescriptor(x2)
    c.PushHistory()
    h = tmp_path / "history.json"
    c.WriteHistory(h)
    assert h.read_text() == "[" + s0 + "," + c.scene_json() + "]"
    assert len(json.loads(h.read_text())) == 2
    c.PopHistory()
    assert c.GetHistorySize() == 1 and c.scene_json() == s0 and np.array_equal(c.GetDescriptor(), x)
    c.PopHistory()
    assert c.GetHistorySize() == 1
    c.ClearHistory()
    assert c.GetHistorySize() == 0
    with pytest.raises(EngineError):
        c.CalculateTorques(3.0)       # the torque query itself is GPU work: no fallback


@pytest.mark.parametrize("name", SCENES)
def test_write_json_is_byte_for_byte_what_json11_writes(name):
    """Core::WriteJson / ConvertCurrentSceneToJson (core_io.cpp:165-169,216-226) against goldens written by the reference's OWN json11
    (oracle/json11_writer.cpp restates core_io.cpp's Json construction over external/json11/json11.cpp, compiled where it lies;
    tests/golden/json11_scene_*.txt). With awkward doubles too, through the helper's canonical form when it is built."""
    import subprocess
    c = Core(scene_path(name))
    golden = open(os.path.join(ROOT, "tests", "golden", f"json11_scene_{name}.txt")).read()
    assert c.scene_json() == golden
    helper = os.path.join(ROOT, "oracle", "_ref", "json11_writer")
    if os.path.exists(helper):
        x = c.GetDescriptor()
        rng = np.random.default_rng(4)
        c.SetDescriptor(x * (1.0 + 1e-3 * rng.standard_normal(x.size)) + 1e-7 * rng.standard_normal(x.size) + np.pi * 1e5 * (rng.random(x.size) < 0.05))
        text = c.scene_json()
        canon = subprocess.run([helper, "canon"], input=text.encode(), capture_output=True, check=True).stdout.decode()
        assert text == canon


def test_generator_chooses_the_curve_group_form_per_scene(tmp_path, monkeypatch):
    """omo_jit.hpp on the host (NVRTC needs no GPU): a scene on sorted curve work lists keeps curve_group<REM> (both forms, chosen at
    launch time), a many-group scene gets curve_group_n<3, REM> (the per-sample form alone); OMO_CURVE_K / OMO_SINCOS_CALL select the
    measured-and-not-kept forms, which still compile."""
    import re

    for k in ("OMO_CURVE_K", "OMO_SINCOS_CALL", "OMO_SORTED_CURVES", "OMO_PLAN", "OMO_JIT_DEFINES"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("OMO_JIT_CACHE", "0")   # compile, do not read or write the in-tree cubin cache

    def forms(scene, tag, **kw):
        prefix = str(tmp_path / tag)
        ok, log = Core(scene).probe_specialised_kernel(dump_prefix=prefix, **kw)
        if not ok and "NVRTC unavailable" in log:
            pytest.skip("no libnvrtc on this machine")
        assert ok, log[:2000]
        src = open(prefix + ".cu").read()
        return set(re.findall(r"w\.(curve_group(?:_n)?<[^>]*>)", src)), src

    stick, src = forms(load_scene("stick-figure"), "stick")
    assert stick and all(f.startswith("curve_group<") for f in stick)
    assert "#define OMO_PLAN 0" in src and "#define OMO_SINCOS_CALL 0" in src
    tree = synthetic.tree_scene(n_joints=24, n_frames=120)
    t3, src = forms(tree, "tree")
    assert t3 and all(f.startswith("curve_group_n<3,") for f in t3)
    assert "#define OMO_PLAN 1" in src
    monkeypatch.setenv("OMO_CURVE_K", "4")
    monkeypatch.setenv("OMO_SINCOS_CALL", "1")
    t4, src = forms(tree, "tree4")
    assert t4 and all(f.startswith("curve_group_n<4,") for f in t4) and "#define OMO_SINCOS_CALL 1" in src
    monkeypatch.setenv("OMO_CURVE_K", "0")
    t0, _ = forms(tree, "tree0")
    assert t0 and all(f.startswith("curve_group<") for f in t0)

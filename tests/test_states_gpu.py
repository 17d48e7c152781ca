"""Next rows N2 / N3 on the GPU: Core::CalculateTorques(t) / SetCurrentState for many t in one launch,
per-frame Euler angles, ExportMotion / history files -- against the goldens of the reference-linked oracle
(tests/golden/states_at.npz) and the CPU oracle. Tolerances as in test_parity_gpu.py."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, SCENES, load_scene, rel_err, scene_path

from optimo_b200.core import Core

pytestmark = pytest.mark.gpu

TIGHT = 2e-11


def _tol(name):
    return 1e-9 if name == "conductor" else TIGHT   # IK angles are converged to 1e-6 only; ulp noise is amplified


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("tag", ["orig", "pert"])
def test_single_t_torques_match_goldens(name, tag):
    g = np.load(os.path.join(GOLDEN, "states_at.npz"))
    c = Core(scene_path(name))
    x, t = g[f"{name}/{tag}/x"], g[f"{name}/{tag}/times"]
    c.SetDescriptor(x)
    tau, q, qd, qdd = c.CalculateTorques(t, want_states=True)      # unsorted, duplicated, closer than 2 dt: several passes
    for k, v in (("tau", tau), ("q", q), ("qd", qd), ("qdd", qdd)):
        assert v.shape == g[f"{name}/{tag}/{k}"].shape
        assert rel_err(v, g[f"{name}/{tag}/{k}"]) <= _tol(name), (k, rel_err(v, g[f"{name}/{tag}/{k}"]))
    # one t at a time, as the reference's UI thread calls it (mainwidget.cpp:190), gives the same rows
    for i in (0, 7, t.size - 6, t.size - 1):
        one = c.CalculateTorques(float(t[i]))
        assert one.shape == (3 * c.J,)
        assert np.array_equal(one, tau[i])
    assert rel_err(c.CalculateEulerAngles(t), g[f"{name}/{tag}/euler"]) <= _tol(name)


@pytest.mark.parametrize("name", SCENES)
def test_objective_window_survives_a_torque_query(name):
    """CalculateTorques borrows the engine's sample grid; the next objective evaluation must see its own window again."""
    g = np.load(os.path.join(GOLDEN, "states_at.npz"))
    c = Core(scene_path(name))
    c.UpdateGlobalWeight(12.5)
    x, x0 = g[f"{name}/pert/x"], g[f"{name}/orig/x"]
    lo, hi = c.window
    f1, t1 = c.CalculateFullObjective(lo, hi, x, x0, want_terms=True)
    c.CalculateTorques(np.arange(c.min_frame_, c.max_frame_ + 1, dtype=np.float64))
    f2, t2 = c.CalculateFullObjective(lo, hi, x, x0, want_terms=True)
    assert f1 == f2 and np.array_equal(t1, t2)
    mt, mc = c.UpdateMaximumTorqueAndCost()
    cache = c.GetOptimizationCache()
    assert abs(mt - np.linalg.norm(cache["tau"], axis=1).max()) <= 1e-12 * mt
    assert mc == cache["local_cost"].max()


def test_single_t_torques_against_the_oracle_on_a_dense_time_list(oracle_port):
    """A heat strip's worth of times (one per pixel column, graphwidget.cpp:325-340): 1500 times closer than 2 dt."""
    name = "stick-figure"
    c = Core(scene_path(name))
    o = oracle_port.Oracle(load_scene(name), "port")
    x = c.GetDescriptor()
    t = np.linspace(c.min_frame_ - 2.0, c.max_frame_ + 2.0, 1500)
    tau = c.CalculateTorques(t)
    ref = o.torques_at(x, t)["tau"]
    assert rel_err(tau, ref) <= TIGHT


@pytest.mark.parametrize("name", SCENES)
def test_export_motion_and_history_files(name, tmp_path, oracle_port):
    c = Core(scene_path(name))
    o = oracle_port.Oracle(load_scene(name), "port")
    x = c.GetDescriptor()
    p = tmp_path / "motion.json"
    c.ExportMotion(p)
    frames = json.loads(p.read_text())
    assert [f["frame"] for f in frames] == list(range(c.min_frame_, c.max_frame_ + 1))
    names = [j["name"] for j in frames[0]["joints"]]
    assert len(names) == c.J
    got = np.array([[v for j in f["joints"] for v in (j["r_x"], j["r_y"], j["r_z"])] for f in frames])
    ref = o.euler_angles_at(x, np.arange(c.min_frame_, c.max_frame_ + 1, dtype=np.float64))
    assert rel_err(got, ref) <= _tol(name)
    # the file is the frame objects joined by ", " inside [ ] (core_io.cpp:252-264), each a json11-style dump
    assert p.read_text() == c.motion_json() and p.read_text().startswith('[{"frame": %d, "joints": [{"name": ' % c.min_frame_)

    # undo stack (core_io.cpp:171-213)
    c.ClearHistory()
    c.PushHistory()
    s0 = c.scene_json()
    x2 = x.copy()
    x2[2::5] += 0.1
    c.SetDescriptor(x2)
    c.PushHistory()
    assert c.GetHistorySize() == 2
    h = tmp_path / "history.json"
    c.WriteHistory(h)
    stamps = json.loads(h.read_text())
    assert len(stamps) == 2 and json.dumps(stamps[0]) != json.dumps(stamps[1])
    assert h.read_text() == "[" + s0 + "," + c.scene_json() + "]"
    c.PopHistory()
    assert c.GetHistorySize() == 1 and c.scene_json() == s0
    assert np.array_equal(c.GetDescriptor(), x)
    c.PopHistory()                                   # "Cannot undo any more."
    assert c.GetHistorySize() == 1
    # the engine follows the undone scene
    tau = c.CalculateTorques(10.0)
    assert rel_err(tau, o.torques_at(x, [10.0])["tau"][0]) <= _tol(name)


@pytest.mark.parametrize("name", SCENES)
def test_export_motion_text_is_json11_canonical(name):
    """Core::ExportMotion (core_io.cpp:252-294) writes json11's dump of every frame. The product's text, parsed and dumped again by
    the reference's own json11 (oracle/_ref/json11_writer canon), must come back byte for byte: key order, separators, %.17g."""
    import os
    import subprocess
    from conftest import ROOT
    helper = os.path.join(ROOT, "oracle", "_ref", "json11_writer")
    if not os.path.exists(helper):
        pytest.skip("oracle/_ref/json11_writer not built (make -C oracle ref, needs /root/reference)")
    c = Core(scene_path(name))
    text = c.motion_json()
    canon = subprocess.run([helper, "canon"], input=text.encode(), capture_output=True, check=True).stdout.decode()
    assert text == canon and text.startswith('[{"frame": ')

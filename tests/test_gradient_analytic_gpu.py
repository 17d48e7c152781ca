"""Next row N4: the analytic gradient (omo_eval_grad_analytic) against the central-difference gradient
of the same objective on the GPU and against the oracle's ("oracle = the existing FD gradient at 1e-6 relative").
The arithmetic itself is checked on the CPU in tests/test_adjoint_math.py."""
import numpy as np
import pytest

from conftest import load_scene, scene_path

from optimo_b200 import synthetic
from optimo_b200.core import Core

pytestmark = pytest.mark.gpu

REL = 1e-6     # of the largest gradient entry: the central difference at h = 1e-4 is itself only that accurate (h^2 f''' / 6)


def _scene(name):
    if name == "stick240":
        return synthetic.stick_240(load_scene("stick-figure"))
    if name == "tree24":
        return synthetic.tree_scene(n_joints=24, n_frames=60)
    return load_scene(name)


@pytest.mark.parametrize("name,w_traj,kernels,cyclic", [
    ("stick-figure", 0.0, [], False), ("stick-figure", 12.5, [(1.0, 20.0, 2.6)], False), ("stick-figure", 12.5, [], True),
    ("fox-tail", 0.0, [], False), ("fox-tail", 3.0, [(1.0, 10.0, 1.0)], True), ("stick240", 12.5, [], False), ("tree24", 1.0, [], False)])
def test_analytic_gradient_matches_the_finite_difference_gradient(oracle_port, name, w_traj, kernels, cyclic):
    scene = _scene(name)
    c = Core(scene)
    c.UpdateGlobalWeight(w_traj); c.set_kernels(kernels); c.set_cyclic(cyclic)
    x0 = c.GetDescriptor()
    x = synthetic.perturbed_candidates(x0, 1, first=5)[0]
    lo, hi = c.window
    f_an, g_an, used = c.CalculateAnalyticGradient(lo, hi, x, x0)
    assert used
    f_fd, g_fd, _ = c.CalculateGradient(lo, hi, x, x0, h=1e-4)
    assert abs(f_an - f_fd) <= 1e-12 * abs(f_fd)            # the per-sample-output kernel (table-driven) vs the specialised one
    scale = max(float(np.max(np.abs(g_fd))), 1.0)
    err = float(np.max(np.abs(g_an - g_fd)))
    assert err <= REL * scale + 2e-14 * abs(f_fd) / 2e-4, (err, scale)
    # every free parameter kind is covered: theta, t_b, t_f entries all carry signal somewhere
    idx = c.free_indices()
    for field in (2, 3, 4):
        sel = idx % 5 == field
        assert np.max(np.abs(g_an[sel])) > 0
    if name in ("stick-figure", "fox-tail"):
        o = oracle_port.Oracle(scene, "port")
        o.set_weight_trajectory(w_traj); o.set_kernels(kernels); o.set_cyclic(cyclic)
        o.update_original(x0)
        fo, go, _ = o.gradient(x, h=1e-4)
        assert np.max(np.abs(g_an - go)) <= REL * max(float(np.max(np.abs(go))), 1.0) + 2e-14 * abs(fo) / 2e-4


def test_analytic_gradient_with_free_times_and_values():
    """t and v of the control points freed as well (the reference's default freezes them, core.cpp:83-88): the derivative with
    respect to a key time goes through the implicit root of BOTH adjacent segments. f is only piecewise smooth in a key time -- a
    key that crosses a sample time switches that sample's segment -- and the shipped keys sit on whole frames, i.e. exactly ON
    sample times, where a central difference straddles the kink. The keys are therefore moved 0.05 frames off the sample grid
    (samples are 0.2 apart, h = 1e-4): there the difference quotient is a fair reference for every field."""
    c = Core(scene_path("stick-figure"))
    x0 = c.GetDescriptor()
    c.freezed_dofs_ = np.zeros(x0.size, dtype=np.int32)
    x = synthetic.perturbed_candidates(x0, 1, first=2)[0]
    x[0::5] += 0.05
    lo, hi = c.window
    f_an, g_an, used = c.CalculateAnalyticGradient(lo, hi, x, x0)
    f_fd, g_fd, _ = c.CalculateGradient(lo, hi, x, x0, h=1e-4)
    assert used and g_an.size == x0.size
    scale = max(float(np.max(np.abs(g_fd))), 1.0)
    assert np.max(np.abs(g_an - g_fd)) <= 2e-6 * scale, np.max(np.abs(g_an - g_fd)) / scale
    field = np.arange(x0.size) % 5
    for k in range(5):
        assert np.max(np.abs(g_an[field == k])) > 0


def test_scenes_outside_the_analytic_gradient_fall_back_to_finite_differences():
    """IK handles (the chain's angles come out of an iterative solve) and an x whose key times are out of order (list-order segment
    search) are answered by the finite-difference gradient, and the caller is told so."""
    c = Core(scene_path("conductor"))
    x0 = c.GetDescriptor()
    lo, hi = c.window
    f, g, used = c.CalculateAnalyticGradient(lo, hi, x0, x0)
    f2, g2, _ = c.CalculateGradient(lo, hi, x0, x0)
    assert not used and f == f2 and np.array_equal(g, g2)
    s = Core(scene_path("stick-figure"))
    xs = s.GetDescriptor()
    x = xs.copy(); x[0] = x[5] + 0.25
    f, g, used = s.CalculateAnalyticGradient(*s.window, x, xs)
    f2, g2, _ = s.CalculateGradient(*s.window, x, xs)
    assert not used and f == f2 and np.array_equal(g, g2)


@pytest.mark.parametrize("name,w_traj", [("stick-figure", 0.0), ("stick-figure", 12.5), ("fox-tail", 12.5)])
def test_reverse_mode_kernel_equals_the_forward_mode_kernel(name, w_traj, monkeypatch):
    """The default (reverse mode: one thread per sample, O(1) passes over the tree) against OMO_ADJ_MODE=forward (one dual-number pass per
    channel and kind): the same gradient to rounding -- they share every kernel behind the per-sample partials."""
    from conftest import scene_path
    from optimo_b200.core import Core
    grads = {}
    for mode in ("reverse", "forward"):
        monkeypatch.setenv("OMO_ADJ_MODE", mode)
        c = Core(scene_path(name))
        c.UpdateGlobalWeight(w_traj)
        x0 = c.GetDescriptor()
        rng = np.random.default_rng(5)
        x = x0.copy(); x[2::5] += 0.3 * (rng.random(x.size // 5) - 0.5)
        lo, hi = c.window
        f, g, used = c.CalculateAnalyticGradient(lo, hi, x, x0)
        assert used
        grads[mode] = (f, g)
    assert grads["reverse"][0] == grads["forward"][0]
    gr, gf = grads["reverse"][1], grads["forward"][1]
    assert np.max(np.abs(gr - gf)) <= 1e-10 * np.max(np.abs(gf)), float(np.max(np.abs(gr - gf)) / np.max(np.abs(gf)))

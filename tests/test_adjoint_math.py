"""CPU check of the analytic gradient's arithmetic (next row N4, optimo_b200/csrc/omo_adjoint.cuh).

The header is generic in its scalar type and compiles for the host; tests/native/adjoint_check.cpp wraps it. Checked here,
without a GPU: the primal restatement against the oracle's per-sample arrays, the forward-mode tangents against central
differences of that primal, the Bezier partials (implicit derivative of the Newton root) against differences of an exactly
solved curve, and the transposed stencil coefficients against the reference's difference formulas. The CUDA kernels that run
the same source are compared with the finite-difference gradient on the GPU in tests/test_gradient_analytic_gpu.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, load_scene, scene_path

from optimo_b200.core import Core

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


@pytest.fixture(scope="module")
def adj():
    so = os.path.join(ROOT, "tests", "native", "_build", "libadjoint_check.so")
    src = os.path.join(ROOT, "tests", "native", "adjoint_check.cpp")
    hdr = os.path.join(ROOT, "optimo_b200", "csrc", "omo_adjoint.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", os.path.join(ROOT, "optimo_b200", "csrc"), "-o", so, src])
    lib = C.CDLL(so)
    lib.adj_sizeof_scene.restype = C.c_ulonglong
    lib.adj_sample_cost.restype = C.c_double
    lib.adj_sample_cost.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, _dp, C.c_longlong, _dp, C.c_int, _dp]
    lib.adj_sample_tangent.restype = C.c_double
    lib.adj_sample_tangent.argtypes = [C.c_void_p, _ip, C.c_int, C.c_int, C.c_int, _dp, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, _dp, C.c_longlong,
                                       _dp, C.c_int]
    lib.adj_sample_reverse.restype = None
    lib.adj_sample_reverse.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, _dp, C.c_longlong, _dp, C.c_int, _dp, _dp, _dp]
    lib.adj_bezier.restype = C.c_double
    lib.adj_bezier.argtypes = [_dp, _dp, C.c_double, C.c_double, C.c_double, _dp]
    lib.adj_d1.restype = C.c_double
    lib.adj_d1.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
    lib.adj_d2.restype = C.c_double
    lib.adj_d2.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
    return lib


def _p(a):
    return a.ctypes.data_as(_dp)


def _keyed_channels(scene):
    """(joint index, axis) of every keyed joint channel, joints in DFS pre-order."""
    out, idx = [], [0]

    def walk(j):
        k = idx[0]; idx[0] += 1
        for a, key in enumerate(("rot_x", "rot_y", "rot_z")):
            if j[key]["curve"] and k > 0:
                out.append((k, a))
        for c in j.get("children", []):
            walk(c)
    walk(scene["joints"])
    return out


@pytest.mark.parametrize("name", ["stick-figure", "fox-tail"])
def test_primal_restatement_matches_the_oracle_and_tangents_match_differences(adj, oracle_port, name):
    scene = load_scene(name)
    core = Core(scene_path(name))
    size = int(adj.adj_sizeof_scene())
    with pytest.raises(Exception):
        core.debug_scene_table(size - 8)                  # the CUDA build's table is not smaller than the host build's ...
    table, sub_end = core.debug_scene_table(size)         # ... and fits it: same layout on both sides
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1)
    x0 = o.descriptor()
    rng = np.random.default_rng(3)
    x = x0.copy(); x[2::5] += 0.3 * (rng.random(x.size // 5) - 0.5)
    o.full_objective(x, x0)
    cache = o.cache()
    J = core.J
    root_e = np.zeros(3)
    # (a) the primal restatement reproduces the oracle's torques and local costs sample by sample
    for i in (0, 1, 2, 57, cache["time"].size - 2, cache["time"].size - 1):
        q, qd, qdd = (np.ascontiguousarray(cache[k][i]) for k in ("q", "qd", "qdd"))
        tau = np.zeros(3 * J)
        interior = int(2 <= i <= cache["time"].size - 3)
        c = adj.adj_sample_cost(table, _p(q), _p(qd), _p(qdd), 1.0, 0.0, 0.2, None, 0, _p(root_e), interior, _p(tau))
        assert np.allclose(tau[3:], cache["tau"][i][3:], rtol=1e-11, atol=1e-12 * np.abs(cache["tau"][i]).max())
        assert abs(c - cache["local_cost"][i]) <= 1e-11 * abs(cache["local_cost"][i]) + 1e-18
    # (b) forward-mode tangents == central differences of the primal, every keyed channel, every kind, with the trajectory term
    i = 57
    q, qd, qdd = (np.ascontiguousarray(cache[k][i]).copy() for k in ("q", "qd", "qdd"))
    p0 = rng.standard_normal(3 * J)
    tau = np.zeros(3 * J)
    w_cost, w_traj, dt = 0.7, 12.5, 0.2
    adj.adj_sample_cost(table, _p(q), _p(qd), _p(qdd), w_cost, w_traj, dt, _p(p0), 1, _p(root_e), 1, _p(tau))
    worst = 0.0
    for (j, axis) in _keyed_channels(scene):
        for kind, arr in enumerate((q, qd, qdd)):
            got = adj.adj_sample_tangent(table, sub_end.ctypes.data_as(_ip), j, axis, kind, _p(q), _p(qd), _p(qdd), _p(tau), w_cost, w_traj, dt, _p(p0), 1,
                                         _p(root_e), 1)
            h = 1e-5 * max(1.0, abs(arr[3 * j + axis]))
            keep = arr[3 * j + axis]
            arr[3 * j + axis] = keep + h
            cp = adj.adj_sample_cost(table, _p(q), _p(qd), _p(qdd), w_cost, w_traj, dt, _p(p0), 1, _p(root_e), 1, None)
            arr[3 * j + axis] = keep - h
            cm = adj.adj_sample_cost(table, _p(q), _p(qd), _p(qdd), w_cost, w_traj, dt, _p(p0), 1, _p(root_e), 1, None)
            arr[3 * j + axis] = keep
            fd = (cp - cm) / (2 * h)
            # central difference: truncation ~h^2, round-off ~eps |c| / h -- both far below 1e-6 |fd| + 1e-9 |c|
            err = abs(got - fd) / (1e-6 * abs(fd) + 1e-9 * abs(cp))
            worst = max(worst, err)
            assert err <= 1.0, (j, axis, kind, got, fd)
    assert worst > 0.0


def _exact_value(ps, pe, ths, the, t):
    p0, p1, p2, p3 = ps[0], ps[0] + ps[4], pe[0] + pe[3], pe[0]
    v0, v1, v2, v3 = ps[1], ps[1] + ps[4] * np.tan(ths), pe[1] + pe[3] * np.tan(the), pe[1]
    a, b, c, d = -p0 + 3 * p1 - 3 * p2 + p3, 3 * (p0 - 2 * p1 + p2), -3 * p0 + 3 * p1, p0 - t
    u = 0.5
    for _ in range(100):
        f = u * (u * (a * u + b) + c) + d
        u -= f / (u * (3 * a * u + 2 * b) + c)
    w = 1 - u
    return v0 * w ** 3 + 3 * v1 * w * w * u + 3 * v2 * w * u * u + v3 * u ** 3


def test_bezier_partials_match_differences_of_the_exactly_solved_curve(adj):
    rng = np.random.default_rng(5)
    for _ in range(50):
        ps = np.array([4.0, rng.normal(), 0.0, -(0.5 + rng.random()), 0.5 + rng.random()])
        pe = np.array([12.0 + rng.random(), rng.normal(), 0.0, -(0.5 + rng.random()), 0.5 + rng.random()])
        ths, the = 0.6 * (rng.random() - 0.5), 0.6 * (rng.random() - 0.5)
        t = ps[0] + (pe[0] - ps[0]) * (0.05 + 0.9 * rng.random())
        out = np.zeros(10)
        val = adj.adj_bezier(_p(ps), _p(pe), ths, the, t, _p(out))
        assert abs(val - _exact_value(ps, pe, ths, the, t)) <= 1e-8        # the reference's Newton stops at |f| < 1e-9
        h = 1e-6
        for k in range(5):
            for which, base in ((0, ps), (1, pe)):
                if k == 2:
                    fd = ((_exact_value(ps, pe, ths + h, the, t) - _exact_value(ps, pe, ths - h, the, t)) if which == 0 else
                          (_exact_value(ps, pe, ths, the + h, t) - _exact_value(ps, pe, ths, the - h, t))) / (2 * h)
                else:
                    up, dn = base.copy(), base.copy()
                    up[k] += h; dn[k] -= h
                    fd = ((_exact_value(up, pe, ths, the, t) - _exact_value(dn, pe, ths, the, t)) if which == 0 else
                          (_exact_value(ps, up, ths, the, t) - _exact_value(ps, dn, ths, the, t))) / (2 * h)
                assert abs(out[5 * which + k] - fd) <= 2e-7 * max(1.0, abs(fd)), (which, k, out[5 * which + k], fd)


def _reference_stencils(q, dt):
    """CalculateFirstCentral / CalculateSecondCentral, OptiMo/core_rbdl.cpp:28-95 (SECOND_ORDER_ACCURATE off)."""
    n = q.size
    d1, d2 = np.zeros(n), np.zeros(n)
    for i in range(n):
        if 2 <= i <= n - 3:
            d1[i] = (-q[i + 2] + 8 * q[i + 1] - 8 * q[i - 1] + q[i - 2]) / (12 * dt)
            d2[i] = (-q[i + 2] + 16 * q[i + 1] - 30 * q[i] + 16 * q[i - 1] - q[i - 2]) / (12 * dt * dt)
        elif i == 0:
            d1[i] = (q[1] - q[0]) / dt
            d2[i] = (q[2] - 2 * q[1] + q[0]) / (2 * dt * dt)
        elif i == n - 1:
            d1[i] = (q[n - 1] - q[n - 2]) / dt
            d2[i] = (q[n - 1] - 2 * q[n - 2] + q[n - 3]) / (2 * dt * dt)
        else:
            d1[i] = (q[i + 1] - q[i - 1]) / (2 * dt)
            d2[i] = (q[i + 1] - 2 * q[i] + q[i - 1]) / (dt * dt)
    return d1, d2


def test_stencil_coefficients_are_the_reference_difference_operators(adj):
    rng = np.random.default_rng(9)
    for n in (3, 4, 5, 6, 11):
        D1 = np.array([[adj.adj_d1(m, i, n, 0.2) for i in range(n)] for m in range(n)])
        D2 = np.array([[adj.adj_d2(m, i, n, 0.2) for i in range(n)] for m in range(n)])
        q = rng.standard_normal(n)
        d1, d2 = _reference_stencils(q, 0.2)
        assert np.allclose(D1 @ q, d1, rtol=1e-13, atol=1e-13) and np.allclose(D2 @ q, d2, rtol=1e-13, atol=1e-12), n


@pytest.mark.parametrize("name", ["stick-figure", "fox-tail", "conductor", "tree16"])
@pytest.mark.parametrize("w_traj", [0.0, 12.5])
def test_reverse_mode_equals_the_forward_mode_tangents_on_every_joint_channel(adj, name, w_traj):
    """omo_ad::sample_reverse (one forward sweep, one fold-back and their hand-written adjoints: every d cost / d (q, qd, qdd) of a
    sample at once) against omo_ad::sample_tangent (one dual-number pass per channel and kind, itself checked against central
    differences above): all 3 x 3(J-1) partials, random states, with and without the trajectory term, chains and trees."""
    from optimo_b200 import synthetic
    scene = synthetic.tree_scene(16, 60) if name == "tree16" else load_scene(name)
    core = Core(scene)
    table, sub_end = core.debug_scene_table(int(adj.adj_sizeof_scene()))
    J = core.J
    rng = np.random.default_rng(11)
    q, qd, qdd = 0.7 * rng.standard_normal(3 * J), rng.standard_normal(3 * J), 3.0 * rng.standard_normal(3 * J)
    p0, root_e = rng.standard_normal(3 * J), 0.3 * rng.standard_normal(3)
    tau = np.zeros(3 * J)
    adj.adj_sample_cost(table, _p(q), _p(qd), _p(qdd), 0.7, w_traj, 0.2, _p(p0), 1, _p(root_e), 1, _p(tau))
    g = [np.zeros(3 * J) for _ in range(3)]
    adj.adj_sample_reverse(table, _p(q), _p(qd), _p(qdd), 0.7, w_traj, 0.2, _p(p0), 1, _p(root_e), 1, _p(g[0]), _p(g[1]), _p(g[2]))
    scale = max(np.abs(np.concatenate(g)).max(), 1e-300)
    for j in range(1, J):
        for axis in range(3):
            for kind in range(3):
                t = adj.adj_sample_tangent(table, sub_end.ctypes.data_as(_ip), j, axis, kind, _p(q), _p(qd), _p(qdd), _p(tau), 0.7, w_traj, 0.2, _p(p0), 1,
                                           _p(root_e), 1)
                assert abs(g[kind][3 * j + axis] - t) <= 1e-11 * abs(t) + 1e-13 * scale, (j, axis, kind, g[kind][3 * j + axis], t)

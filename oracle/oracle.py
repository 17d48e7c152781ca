"""TEST INFRASTRUCTURE -- ctypes front end of the CPU oracle (oracle/oracle.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module. The product package (optimo_b200/) never does.

`Oracle(kind="port")` loads oracle/liboracle.so (restatement); `Oracle(kind="ref")` loads
oracle/_ref/liboracle_ref.so (same glue over the reference's own RBDL/curve objects, prebuilt in the
authoring container because /root/reference does not exist on the GPU box).

Scene JSON is parsed here with Python's json module and replayed through the oracle's builder calls in
the order OptiMo/core_io.cpp:112-151,228-250 constructs the object graph.
"""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATHS = {"port": os.path.join(_HERE, "liboracle.so"), "ref": os.path.join(_HERE, "_ref", "liboracle_ref.so")}
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def available(kind):
    return os.path.exists(_PATHS[kind])


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_dp)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_ip)


_LIBS = {}


def _load(kind):
    if kind in _LIBS:
        return _LIBS[kind]
    lib = C.CDLL(_PATHS[kind])
    vp, cd, ci, cs = C.c_void_p, C.c_double, C.c_int, C.c_char_p
    sig = {
        "orc_backend": (cs, []),
        "orc_scene_new": (vp, []),
        "orc_scene_free": (None, [vp]),
        "orc_add_joint": (ci, [vp, ci, _dp, _dp, cs]),
        "orc_joint_add_cp": (None, [vp, ci, ci, cd, cd, cd, cd, cd]),
        "orc_add_ik": (ci, [vp, cs, ci, ci, _dp]),
        "orc_ik_add_cp": (None, [vp, ci, ci, cd, cd, cd, cd, cd]),
        "orc_finalize": (None, [vp, ci, ci]),
        "orc_n_joints": (ci, [vp]),
        "orc_n_items": (ci, [vp]),
        "orc_descriptor_size": (ci, [vp]),
        "orc_get_descriptor": (None, [vp, _dp]),
        "orc_set_descriptor": (None, [vp, _dp, ci]),
        "orc_has_degenerate_inertia": (ci, [vp]),
        "orc_set_cyclic": (None, [vp, ci]),
        "orc_set_weight_trajectory": (None, [vp, cd]),
        "orc_set_per_item_weight": (None, [vp, ci, cd]),
        "orc_clear_kernels": (None, [vp]),
        "orc_add_kernel": (None, [vp, cd, cd, cd]),
        "orc_set_sampling_concurrency": (None, [vp, ci]),
        "orc_update_original": (None, [vp, _dp, ci]),
        "orc_get_original": (None, [vp, _dp]),
        "orc_set_selected": (None, [vp, ci, ci]),
        "orc_update_freezed_from_selection": (None, [vp]),
        "orc_set_freezed": (None, [vp, _ip, ci]),
        "orc_get_freezed": (None, [vp, _ip]),
        "orc_cost_weight": (cd, [vp, cd]),
        "orc_parameter_weights": (None, [vp, _dp]),
        "orc_time_grid": (ci, [cd, cd, cd, _dp, ci]),
        "orc_full_objective": (cd, [vp, cd, cd, _dp, _dp, ci, _dp]),
        "orc_cache_samples": (ci, [vp]),
        "orc_dof": (ci, [vp]),
        "orc_get_cache": (None, [vp, ci, _dp]),
        "orc_positions": (ci, [vp, cd, cd, _dp, ci, _dp]),
        "orc_torques_at": (None, [vp, _dp, ci, _dp, ci, _dp, _dp, _dp, _dp]),
        "orc_euler_angles_at": (None, [vp, _dp, ci, _dp, ci, _dp]),
        "orc_evaluate_objective": (cd, [vp, _dp, ci, _dp, ci, cd, cd]),
        "orc_gradient": (cd, [vp, _dp, ci, cd, cd, cd, _dp, _dp]),
        "orc_probe_inverse_dynamics": (None, [ci, _ip, _ip, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
        "orc_probe_inverse_kinematics": (ci, [ci, _dp, _dp, _dp, _dp, _dp, _dp, cd, cd, ci, _dp, _ip]),
        "orc_probe_curve": (None, [ci, _dp, ci, _dp, _dp]),
        "orc_var_chain": (ci, [vp, ci, ci, _ip]),
        "orc_joint_physical": (None, [vp, ci, _dp]),
        "orc_counters": (None, [C.POINTER(C.c_long), ci]),
        "orc_hardware_concurrency": (ci, []),
        "orc_bench_batch": (cd, [vp, _dp, ci, ci, cd, cd, ci, ci, _dp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _LIBS[kind] = lib
    return lib


def time_grid(min_t, max_t, dt=0.2, kind="port"):
    lib = _load(kind)
    n = lib.orc_time_grid(min_t, max_t, dt, None, 0)
    out = np.zeros(n)
    lib.orc_time_grid(min_t, max_t, dt, out.ctypes.data_as(_dp), n)
    return out


class Oracle:
    """One scene loaded into the oracle (== a headless reference `Core` + `Object`)."""

    def __init__(self, scene, kind="port"):
        if isinstance(scene, (str, os.PathLike)):
            with open(scene) as fh:
                scene = json.load(fh)
        self.kind = kind
        self.lib = _load(kind)
        self.h = self.lib.orc_scene_new()
        self.joint_names = []
        self.item_names = []
        self._add_joint(scene["joints"], -1)
        self.item_names = list(self.joint_names)
        for ik in scene.get("ik_handlers", []):
            root, end = self.joint_names.index(ik["root_name"]), self.joint_names.index(ik["end_effector_name"])
            _, vp = _d([ik["pos_x"]["var"], ik["pos_y"]["var"], ik["pos_z"]["var"]])
            k = self.lib.orc_add_ik(self.h, ik["name"].encode(), root, end, vp)
            for axis, key in enumerate(("pos_x", "pos_y", "pos_z")):
                for cp in ik[key]["curve"]:
                    self.lib.orc_ik_add_cp(self.h, k, axis, *[float(v) for v in cp])
            self.item_names.append(ik["name"])
        self.min_frame, self.max_frame = int(scene["min_frame"]), int(scene["max_frame"])
        self.lib.orc_finalize(self.h, self.min_frame, self.max_frame)
        self.D = self.lib.orc_descriptor_size(self.h)
        self.dof = self.lib.orc_dof(self.h)
        self.J = self.lib.orc_n_joints(self.h)

    def _add_joint(self, j, parent):
        _, op = _d(j["offset"])
        _, vp = _d([j["rot_x"]["var"], j["rot_y"]["var"], j["rot_z"]["var"]])
        idx = self.lib.orc_add_joint(self.h, parent, op, vp, j["name"].encode())
        self.joint_names.append(j["name"])
        for axis, key in enumerate(("rot_x", "rot_y", "rot_z")):
            for cp in j[key]["curve"]:
                self.lib.orc_joint_add_cp(self.h, idx, axis, *[float(v) for v in cp])
        for c in j.get("children", []):
            self._add_joint(c, idx)

    def __del__(self):
        try:
            self.lib.orc_scene_free(self.h)
        except Exception:
            pass

    # --- window of the objective (core_opt.cpp:125-126) -------------------------------------------
    @property
    def window(self):
        return self.min_frame - 2.0, self.max_frame + 2.0

    def descriptor(self):
        x = np.zeros(self.D)
        self.lib.orc_get_descriptor(self.h, x.ctypes.data_as(_dp))
        return x

    def original(self):
        x = np.zeros(self.D)
        self.lib.orc_get_original(self.h, x.ctypes.data_as(_dp))
        return x

    def update_original(self, x):
        a, p = _d(x)
        self.lib.orc_update_original(self.h, p, a.size)

    def set_weight_trajectory(self, w):
        self.lib.orc_set_weight_trajectory(self.h, float(w))

    def set_per_item_weight(self, name, w):
        self.lib.orc_set_per_item_weight(self.h, self.item_names.index(name), float(w))

    def set_kernels(self, kernels):
        self.lib.orc_clear_kernels(self.h)
        for amp, mu, sigma in kernels:
            self.lib.orc_add_kernel(self.h, float(amp), float(mu), float(sigma))

    def set_cyclic(self, flag):
        self.lib.orc_set_cyclic(self.h, int(bool(flag)))

    def set_sampling_concurrency(self, c):
        self.lib.orc_set_sampling_concurrency(self.h, int(c))

    def select_items(self, names):
        for k, n in enumerate(self.item_names):
            self.lib.orc_set_selected(self.h, k, int(n in names))
        self.lib.orc_update_freezed_from_selection(self.h)

    def freezed(self):
        f = np.zeros(self.D, dtype=np.int32)
        self.lib.orc_get_freezed(self.h, f.ctypes.data_as(_ip))
        return f

    def set_freezed(self, f):
        a, p = _i(f)
        self.lib.orc_set_freezed(self.h, p, a.size)

    def free_indices(self):
        return np.nonzero(self.freezed() == 0)[0].astype(np.int32)

    def parameter_weights(self):
        w = np.zeros(self.D)
        self.lib.orc_parameter_weights(self.h, w.ctypes.data_as(_dp))
        return w

    def cost_weight(self, t):
        return self.lib.orc_cost_weight(self.h, float(t))

    def full_objective(self, x, x_original=None, window=None):
        """Returns (f, (C, D_param, D_traj))."""
        lo, hi = window or self.window
        xa, xp = _d(x)
        oa, op = _d(self.original() if x_original is None else x_original)
        terms = np.zeros(3)
        f = self.lib.orc_full_objective(self.h, lo, hi, xp, op, xa.size, terms.ctypes.data_as(_dp))
        return f, terms

    def cache(self):
        n = self.lib.orc_cache_samples(self.h)
        out = {}
        for which, (name, width) in enumerate(
            [("time", 0), ("q", self.dof), ("qd", self.dof), ("qdd", self.dof), ("tau", self.dof), ("local_cost", 0)]
        ):
            a = np.zeros((n, width)) if width else np.zeros(n)
            self.lib.orc_get_cache(self.h, which, a.ctypes.data_as(_dp))
            out[name] = a
        return out

    def p_original(self):
        n = self.lib.orc_cache_samples(self.h)
        a = np.zeros((n, 3 * self.J))
        self.lib.orc_get_cache(self.h, 6, a.ctypes.data_as(_dp))
        return a

    def positions(self, x, window=None):
        lo, hi = window or self.window
        n = self.lib.orc_time_grid(lo, hi, 0.2, None, 0)
        xa, xp = _d(x)
        p = np.zeros((n, 3 * self.J))
        self.lib.orc_positions(self.h, lo, hi, xp, xa.size, p.ctypes.data_as(_dp))
        return p

    def torques_at(self, x, times):
        """Core::CalculateTorques(t) (core_rbdl.cpp:264-289) for every t: dict of tau, q, qd, qdd, each [M, dof]."""
        xa, xp = _d(x)
        ta, tp = _d(times)
        out = {k: np.zeros((ta.size, self.dof)) for k in ("tau", "q", "qd", "qdd")}
        self.lib.orc_torques_at(self.h, xp, xa.size, tp, ta.size, *[out[k].ctypes.data_as(_dp) for k in ("tau", "q", "qd", "qdd")])
        return out

    def euler_angles_at(self, x, times):
        """Joint::GetEulerAngle(t) of every joint (root included), [M, 3J] -- what Core::ExportMotion writes."""
        xa, xp = _d(x)
        ta, tp = _d(times)
        e = np.zeros((ta.size, 3 * self.J))
        self.lib.orc_euler_angles_at(self.h, xp, xa.size, tp, ta.size, e.ctypes.data_as(_dp))
        return e

    def gradient(self, x, h=1e-4, window=None):
        """Central-difference gradient on the free parameters. Returns (f0, grad[F], fpm[F,2])."""
        lo, hi = window or self.window
        xa, xp = _d(x)
        F = int((self.freezed() == 0).sum())
        g, fpm = np.zeros(F), np.zeros((F, 2))
        f0 = self.lib.orc_gradient(self.h, xp, xa.size, lo, hi, float(h), g.ctypes.data_as(_dp), fpm.ctypes.data_as(_dp))
        return f0, g, fpm

    def chain(self, item, axis):
        order = np.zeros(4096, dtype=np.int32)
        n = self.lib.orc_var_chain(self.h, item, axis, order.ctypes.data_as(_ip))
        return order[:n].copy()

    def joint_physical(self, joint):
        out = np.zeros(14)
        self.lib.orc_joint_physical(self.h, joint, out.ctypes.data_as(_dp))
        return out[0], out[1:4].copy(), out[4:13].reshape(3, 3).copy()

    def bench_batch(self, X, mode, threads=0, window=None):
        """CPU baseline; returns (seconds, f[B]). mode 0 = ref-as-is, 1 = ref-candidate-parallel."""
        lo, hi = window or self.window
        Xa, Xp = _d(X)
        B, D = Xa.shape
        f = np.zeros(B)
        sec = self.lib.orc_bench_batch(self.h, Xp, B, D, lo, hi, int(mode), int(threads), f.ctypes.data_as(_dp))
        return sec, f


def counters(kind="port", reset=False):
    out = (C.c_long * 4)()
    _load(kind).orc_counters(out, int(reset))
    return {"bezier_evals": out[0], "newton_iters": out[1], "ik_solves": out[2], "ik_iters": out[3]}


def hardware_concurrency(kind="port"):
    return _load(kind).orc_hardware_concurrency()


def probe_inverse_dynamics(kind, parent, jkind, xt, mass, com, inertia, gravity, q, qd, qdd):
    lib = _load(kind)
    n = len(parent)
    tau = np.zeros(3 * n)
    lib.orc_probe_inverse_dynamics(
        n, _i(parent)[1], _i(jkind)[1], _d(xt)[1], _d(mass)[1], _d(com)[1], _d(inertia)[1], _d(gravity)[1], _d(q)[1],
        _d(qd)[1], _d(qdd)[1], tau.ctypes.data_as(_dp))
    return tau


def probe_inverse_kinematics(kind, xt, mass, com, inertia, point, target, tol=1e-6, lam=0.4, max_iter=20):
    lib = _load(kind)
    n = len(mass)
    q = np.zeros(3 * n)
    it = C.c_int(0)
    ok = lib.orc_probe_inverse_kinematics(
        n, _d(xt)[1], _d(mass)[1], _d(com)[1], _d(inertia)[1], _d(point)[1], _d(target)[1], tol, lam, max_iter,
        q.ctypes.data_as(_dp), C.byref(it))
    return bool(ok), q, it.value


def probe_curve(kind, cps, ts):
    lib = _load(kind)
    cps = np.ascontiguousarray(cps, dtype=np.float64).reshape(-1, 5)
    ts = np.ascontiguousarray(ts, dtype=np.float64)
    out = np.zeros(ts.size)
    lib.orc_probe_curve(cps.shape[0], cps.ctypes.data_as(_dp), ts.size, ts.ctypes.data_as(_dp), out.ctypes.data_as(_dp))
    return out

"""ctypes view of the headless Core (include/optimo_core.h) and of the raw engine ABI (include/omo.h).

Method names follow the reference's `Core` (OptiMo/core.h) so that tests read like calls into the reference.
Nothing here computes: every number comes out of liboptimo_core.so / libomo.so (CUDA). If the libraries are
missing, import fails loudly -- there is no CPU path.
"""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_CORE_SO = os.path.join(_HERE, "host", "liboptimo_core.so")
_OMO_SO = os.path.join(_HERE, "csrc", "libomo.so")

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p


class EngineError(RuntimeError):
    pass


class ForcedStop(RuntimeError):
    """Mirror of nlopt::forced_stop thrown by EvaluateObjective (core_opt.cpp:50-54)."""


def lib_paths():
    return {"core": _CORE_SO, "omo": _OMO_SO}


def _load():
    for p in (_OMO_SO, _CORE_SO):
        if not os.path.exists(p):
            raise ImportError(f"{p} is missing: run `make -C optimo_b200` (or __graft_entry__.build()); there is no fallback path")
    omo = C.CDLL(_OMO_SO, mode=C.RTLD_GLOBAL)
    core = C.CDLL(_CORE_SO)
    return omo, core


_omo, _core = _load()


def _sig(lib, name, res, args):
    fn = getattr(lib, name)
    fn.restype = res
    fn.argtypes = args
    return fn


ci, cd, cs = C.c_int, C.c_double, C.c_char_p
_CORE_SIG = {
    "optimo_last_error": (cs, []),
    "optimo_core_new": (_vp, []),
    "optimo_core_instance": (_vp, []),
    "optimo_core_free": (None, [_vp]),
    "optimo_read_json": (ci, [_vp, cs]),
    "optimo_read_json_string": (ci, [_vp, cs]),
    "optimo_write_json": (ci, [_vp, cs]),
    "optimo_scene_json": (ci, [_vp, C.c_char_p, ci]),
    "optimo_set_device": (ci, [_vp, ci]),
    "optimo_set_devices": (ci, [_vp, _ip, ci]),
    "optimo_min_frame": (ci, [_vp]),
    "optimo_max_frame": (ci, [_vp]),
    "optimo_set_frames": (None, [_vp, ci, ci]),
    "optimo_num_joints": (ci, [_vp]),
    "optimo_num_items": (ci, [_vp]),
    "optimo_item_name": (ci, [_vp, ci, C.c_char_p, ci]),
    "optimo_descriptor_size": (ci, [_vp]),
    "optimo_get_descriptor": (ci, [_vp, _dp]),
    "optimo_set_descriptor": (ci, [_vp, _dp, ci]),
    "optimo_set_cyclic": (None, [_vp, ci]),
    "optimo_update_original_parameters": (ci, [_vp, _dp, ci]),
    "optimo_retrieve_original_parameters": (ci, [_vp, _dp]),
    "optimo_update_global_weight": (None, [_vp, cd]),
    "optimo_update_per_item_weight": (ci, [_vp, cd, cs]),
    "optimo_clear_kernels": (None, [_vp]),
    "optimo_add_kernel": (None, [_vp, cd, cd, cd]),
    "optimo_get_cost_weight": (cd, [_vp, cd]),
    "optimo_parameter_weights": (ci, [_vp, _dp]),
    "optimo_set_selected": (ci, [_vp, cs, ci]),
    "optimo_update_freezed_dofs_from_selection": (None, [_vp]),
    "optimo_get_freezed_dofs": (ci, [_vp, _ip]),
    "optimo_set_freezed_dofs": (ci, [_vp, _ip, ci]),
    "optimo_joint_physical": (ci, [_vp, ci, _dp]),
    "optimo_calculate_full_objective": (ci, [_vp, cd, cd, _dp, _dp, ci, _dp, _dp]),
    "optimo_calculate_full_objective_batch": (ci, [_vp, cd, cd, _dp, ci, _dp, ci, _dp, _dp]),
    "optimo_calculate_gradient": (ci, [_vp, cd, cd, _dp, _dp, ci, _ip, ci, cd, _dp, _dp, _dp]),
    "optimo_positions": (ci, [_vp, cd, cd, _dp, ci, _dp]),
    "optimo_num_samples": (ci, [_vp, cd, cd]),
    "optimo_calculate_torques": (ci, [_vp, _dp, ci, _dp, _dp, _dp, _dp]),
    "optimo_euler_angles": (ci, [_vp, _dp, ci, _dp]),
    "optimo_update_maximum_torque_and_cost": (ci, [_vp, _dp, _dp]),
    "optimo_export_motion": (ci, [_vp, cs]),
    "optimo_motion_json": (ci, [_vp, C.c_char_p, ci]),
    "optimo_push_history": (ci, [_vp]),
    "optimo_pop_history": (ci, [_vp]),
    "optimo_clear_history": (None, [_vp]),
    "optimo_history_size": (ci, [_vp]),
    "optimo_write_history": (ci, [_vp, cs]),
    "optimo_evaluate_objective": (ci, [_vp, _dp, ci, _dp, ci, cd, cd, ci, cd, _dp, _dp]),
    "optimo_set_stop_optimization": (None, [_vp, ci]),
    "optimo_get_stop_optimization": (ci, [_vp]),
    "optimo_perform_optimization": (ci, [_vp, ci, _dp, _ip]),
    "optimo_perform_optimization_gradient": (ci, [_vp, ci, cs, _dp, _ip]),
    "optimo_set_batched_start": (None, [_vp, ci]),
    "optimo_last_solve_batched": (ci, [_vp]),
    "optimo_probe_specialised_kernel": (ci, [_vp, ci, ci, cs, cs, ci]),
    "optimo_debug_scene_table": (ci, [_vp, _vp, C.c_ulonglong, _ip]),
    "optimo_calculate_analytic_gradient": (ci, [_vp, cd, cd, _dp, _dp, ci, _ip, ci, _dp, _dp, _ip]),
    "optimo_record_solve_trace": (None, [_vp, ci]),
    "optimo_solve_trace_size": (ci, [_vp, _ip, _ip]),
    "optimo_get_solve_trace": (ci, [_vp, _dp]),
    "optimo_cache_size": (ci, [_vp, _ip, _ip]),
    "optimo_get_cache": (ci, [_vp, ci, _dp]),
    "optimo_clear_optimization_cache": (None, [_vp]),
    "optimo_engine": (_vp, [_vp]),
    "optimo_engine_synced": (_vp, [_vp, cd, cd]),
    "optimo_engine_multi": (_vp, [_vp]),
}
for _n, (_r, _a) in _CORE_SIG.items():
    _sig(_core, _n, _r, _a)

_ull = C.c_ulonglong
_OMO_SIG = {
    "omo_last_error": (cs, [_vp]),
    "omo_version": (cs, []),
    "omo_num_samples": (ci, [_vp]),
    "omo_descriptor_size": (ci, [_vp]),
    "omo_dof": (ci, [_vp]),
    "omo_num_walked_bodies": (ci, [_vp, _ip]),
    "omo_get_time_grid": (ci, [_vp, _dp]),
    "omo_get_cost_weights": (ci, [_vp, _dp]),
    "omo_get_p_original": (ci, [_vp, _dp]),
    "omo_eval": (ci, [_vp, _dp, ci, _dp, _dp]),
    "omo_eval_free": (ci, [_vp, _dp, _ip, ci, _dp, ci, _dp, _dp]),
    "omo_eval_device": (ci, [_vp, _vp, ci, _vp, _vp, ci]),
    "omo_eval_device_static": (ci, [_vp, _vp, ci, _vp, _vp, ci]),
    "omo_set_batch_hint": (ci, [_vp, ci]),
    "omo_peer_gather_create": (ci, [_vp, ci, ci, ci, C.POINTER(_vp), C.c_char_p]),
    "omo_peer_gather_connect": (ci, [_vp, C.c_char_p]),
    "omo_peer_gather_eval": (ci, [_vp, _vp, ci, ci, ci]),
    "omo_peer_gather_values": (_vp, [_vp]),
    "omo_peer_gather_error": (ci, [_vp]),
    "omo_peer_gather_destroy": (None, [_vp]),
    "omo_eval_host_in": (ci, [_vp, _vp, ci, _vp, _vp, ci]),
    "omo_eval_grad": (ci, [_vp, _dp, _ip, ci, cd, _dp, _dp, _dp]),
    "omo_eval_grad_columns": (ci, [_vp, _dp, _ip, ci, cd, ci, ci, _vp, ci]),
    "omo_grad_from_columns": (ci, [_dp, ci, cd, _dp, _dp]),
    "omo_shard_range": (None, [ci, ci, ci, _ip, _ip]),
    "omo_device_alloc": (ci, [_vp, _ull, C.POINTER(_vp)]),
    "omo_device_free": (ci, [_vp, _vp]),
    "omo_host_alloc_pinned": (ci, [_vp, _ull, C.POINTER(_vp)]),
    "omo_host_free_pinned": (ci, [_vp, _vp]),
    "omo_memcpy_h2d": (ci, [_vp, _vp, _vp, _ull]),
    "omo_memcpy_d2h": (ci, [_vp, _vp, _vp, _ull]),
    "omo_synchronize": (ci, [_vp]),
    "omo_stream": (_vp, [_vp]),
    "omo_set_counting": (ci, [_vp, ci]),
    "omo_get_counters": (ci, [_vp, C.POINTER(_ull)]),
    "omo_kernel_timing": (ci, [_vp, ci, _dp, _ip]),
    "omo_measure_dfma_peak": (ci, [_vp, _dp]),
    "omo_jit_info": (ci, [_vp, C.POINTER(_ull), cs, ci]),
    "omo_jit_describe": (ci, [_vp, C.c_char_p, ci]),
    "omo_selftest_division": (ci, [_vp, _dp, _dp, ci, _dp, _dp, _ip]),
    "omo_grad_sparsity": (ci, [_vp, C.POINTER(_ull)]),
    "omo_graph_replays": (_ull, [_vp]),
    "omo_multi_last_error": (cs, [_vp]),
    "omo_multi_num_devices": (ci, [_vp]),
    "omo_multi_device": (ci, [_vp, ci]),
    "omo_multi_ctx": (_vp, [_vp, ci]),
    "omo_multi_set_gather": (ci, [_vp, ci]),
    "omo_multi_get_gather": (ci, [_vp]),
    "omo_multi_set_window": (ci, [_vp, cd, cd, cd]),
    "omo_multi_eval": (ci, [_vp, _dp, ci, _dp, _dp]),
    "omo_multi_eval_free": (ci, [_vp, _dp, _ip, ci, _dp, ci, _dp, _dp]),
    "omo_multi_eval_grad": (ci, [_vp, _dp, _ip, ci, cd, _dp, _dp, _dp]),
}
for _n, (_r, _a) in _OMO_SIG.items():
    _sig(_omo, _n, _r, _a)


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_dp)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_ip)


def shard_range(n, rank, world):
    b, c = C.c_int(0), C.c_int(0)
    _omo.omo_shard_range(n, rank, world, C.byref(b), C.byref(c))
    return b.value, c.value


def grad_from_columns(f_cols, h):
    f_cols, fp = _d(f_cols)
    F = (f_cols.size - 1) // 2
    g = np.zeros(F)
    f0 = C.c_double(0.0)
    _omo.omo_grad_from_columns(fp, F, float(h), C.byref(f0), g.ctypes.data_as(_dp))
    return f0.value, g


class Engine:
    """Raw view of an omo_ctx* (include/omo.h) owned by a Core. Device-resident entry points for the harness."""

    def __init__(self, handle):
        if not handle:
            raise EngineError("engine is not available: " + (_core.optimo_last_error() or b"").decode())
        self.h = handle

    def _ck(self, st, what):
        if st != 0:
            raise EngineError(f"{what}: status {st}: {(_omo.omo_last_error(self.h) or b'').decode()}")

    @property
    def num_samples(self):
        return _omo.omo_num_samples(self.h)

    @property
    def D(self):
        return _omo.omo_descriptor_size(self.h)

    @property
    def dof(self):
        return _omo.omo_dof(self.h)

    def walked_bodies(self):
        """(bodies the batch kernels run per sample-eval, how many of them take the general EulerXYZ step)."""
        g = C.c_int(0)
        n = _omo.omo_num_walked_bodies(self.h, C.byref(g))
        return n, g.value

    def time_grid(self):
        t = np.zeros(self.num_samples)
        self._ck(_omo.omo_get_time_grid(self.h, t.ctypes.data_as(_dp)), "omo_get_time_grid")
        return t

    def cost_weights(self):
        w = np.zeros(self.num_samples)
        self._ck(_omo.omo_get_cost_weights(self.h, w.ctypes.data_as(_dp)), "omo_get_cost_weights")
        return w

    def p_original(self):
        p = np.zeros((self.num_samples, self.dof))
        self._ck(_omo.omo_get_p_original(self.h, p.ctypes.data_as(_dp)), "omo_get_p_original")
        return p

    def eval(self, X, want_terms=False):
        X, xp = _d(X)
        X2 = X.reshape(-1, self.D)
        f = np.zeros(X2.shape[0])
        terms = np.zeros((X2.shape[0], 3)) if want_terms else None
        self._ck(_omo.omo_eval(self.h, xp, X2.shape[0], f.ctypes.data_as(_dp), terms.ctypes.data_as(_dp) if want_terms else None), "omo_eval")
        return (f, terms) if want_terms else f

    def eval_host_ptr(self, x_ptr, B, f_ptr):
        """omo_eval on caller-owned (e.g. pinned) host buffers given as integer addresses."""
        self._ck(_omo.omo_eval(self.h, C.cast(x_ptr, _dp), B, C.cast(f_ptr, _dp), None), "omo_eval")

    def eval_free(self, x_base, free_idx, Xfree, want_terms=False):
        """omo_eval_free: candidates given by their free entries only (the callback's own parameter space)."""
        xb, xbp = _d(x_base)
        idx, ip = _i(free_idx)
        Xf = np.ascontiguousarray(Xfree, dtype=np.float64).reshape(-1, idx.size)
        f = np.zeros(Xf.shape[0])
        terms = np.zeros((Xf.shape[0], 3)) if want_terms else None
        self._ck(_omo.omo_eval_free(self.h, xbp, ip, idx.size, Xf.ctypes.data_as(_dp), Xf.shape[0], f.ctypes.data_as(_dp),
                                    terms.ctypes.data_as(_dp) if want_terms else None), "omo_eval_free")
        return (f, terms) if want_terms else f

    def eval_free_host_ptr(self, x_base, free_idx, xf_ptr, B, f_ptr):
        """omo_eval_free on caller-owned (e.g. pinned) host buffers given as integer addresses."""
        xb, xbp = _d(x_base)
        idx, ip = _i(free_idx)
        self._ck(_omo.omo_eval_free(self.h, xbp, ip, idx.size, C.cast(xf_ptr, _dp), B, C.cast(f_ptr, _dp), None), "omo_eval_free")

    def eval_device(self, dX, B, dF, dTerms=None, sync=True):
        self._ck(_omo.omo_eval_device(self.h, dX, B, dF, dTerms, int(sync)), "omo_eval_device")

    def eval_device_static(self, dX, B, dF, dTerms=None, sync=True):
        """omo_eval_device_static: the same batch with its scratch in the library's static device array (no allocation)."""
        self._ck(_omo.omo_eval_device_static(self.h, dX, B, dF, dTerms, int(sync)), "omo_eval_device_static")

    def set_batch_hint(self, total):
        """This engine evaluates slices of a batch of `total` candidates (kernel choice by the whole batch; 0 = off)."""
        self._ck(_omo.omo_set_batch_hint(self.h, int(total)), "omo_set_batch_hint")

    def eval_host_in(self, x_ptr, B, dF, dTerms=None, sync=False):
        """omo_eval_host_in: host (pinned) candidates in, values left on the device."""
        self._ck(_omo.omo_eval_host_in(self.h, x_ptr, B, dF, dTerms, int(sync)), "omo_eval_host_in")

    def eval_grad(self, x, free_idx, h=1e-4):
        x, xp = _d(x)
        idx, ip = _i(free_idx)
        g, fpm = np.zeros(idx.size), np.zeros((idx.size, 2))
        f0 = C.c_double(0.0)
        self._ck(_omo.omo_eval_grad(self.h, xp, ip, idx.size, float(h), C.byref(f0), g.ctypes.data_as(_dp), fpm.ctypes.data_as(_dp)), "omo_eval_grad")
        return f0.value, g, fpm

    def eval_grad_columns(self, x, free_idx, h, col_begin, col_count, d_out, sync=True):
        x, xp = _d(x)
        idx, ip = _i(free_idx)
        self._ck(_omo.omo_eval_grad_columns(self.h, xp, ip, idx.size, float(h), col_begin, col_count, d_out, int(sync)), "omo_eval_grad_columns")

    def device_alloc(self, nbytes):
        p = _vp()
        self._ck(_omo.omo_device_alloc(self.h, nbytes, C.byref(p)), "omo_device_alloc")
        return p.value

    def device_free(self, p):
        self._ck(_omo.omo_device_free(self.h, p), "omo_device_free")

    def pinned_alloc(self, nbytes):
        p = _vp()
        self._ck(_omo.omo_host_alloc_pinned(self.h, nbytes, C.byref(p)), "omo_host_alloc_pinned")
        return p.value

    def pinned_free(self, p):
        self._ck(_omo.omo_host_free_pinned(self.h, p), "omo_host_free_pinned")

    def pinned_array(self, shape):
        """numpy view over freshly allocated pinned host memory (freed with pinned_free(arr.ctypes.data))."""
        n = int(np.prod(shape))
        p = self.pinned_alloc(n * 8)
        buf = (C.c_double * n).from_address(p)
        return np.frombuffer(buf, dtype=np.float64).reshape(shape)

    def h2d(self, dst, arr):
        arr = np.ascontiguousarray(arr)
        self._ck(_omo.omo_memcpy_h2d(self.h, dst, arr.ctypes.data, arr.nbytes), "omo_memcpy_h2d")

    def d2h(self, src, shape, dtype=np.float64):
        out = np.zeros(shape, dtype=dtype)
        self._ck(_omo.omo_memcpy_d2h(self.h, out.ctypes.data, src, out.nbytes), "omo_memcpy_d2h")
        return out

    def synchronize(self):
        self._ck(_omo.omo_synchronize(self.h), "omo_synchronize")

    def stream(self):
        return _omo.omo_stream(self.h)

    def set_counting(self, on):
        self._ck(_omo.omo_set_counting(self.h, int(on)), "omo_set_counting")

    def counters(self):
        out = (_ull * 5)()
        self._ck(_omo.omo_get_counters(self.h, out), "omo_get_counters")
        return {"bezier_evals": out[0], "newton_iters": out[1], "ik_solves": out[2], "ik_iters": out[3], "launches": out[4]}

    def graph_replays(self):
        """Single-candidate evaluations served by replaying the captured CUDA graph (include/omo.h)."""
        return int(_omo.omo_graph_replays(self.h))

    def kernel_timing(self, reset=0):
        ms, n = C.c_double(0.0), C.c_int(0)
        self._ck(_omo.omo_kernel_timing(self.h, int(reset), C.byref(ms), C.byref(n)), "omo_kernel_timing")
        return ms.value, n.value

    def jit_info(self):
        out = (_ull * 4)()
        buf = C.create_string_buffer(1 << 14)
        self._ck(_omo.omo_jit_info(self.h, out, buf, len(buf)), "omo_jit_info")
        return {"launches": int(out[0]), "instances": int(out[1]), "max_registers": int(out[2]), "compile_ms": int(out[3]),
                "log": buf.value.decode(errors="replace")}

    def jit_describe(self):
        """The specialised kernel instances this engine built (form, registers, shared memory, build time, cache hit)."""
        buf = C.create_string_buffer(1 << 14)
        self._ck(_omo.omo_jit_describe(self.h, buf, len(buf)), "omo_jit_describe")
        return json.loads(buf.value.decode() or "[]")

    def selftest_division(self, num, den):
        """(q_fast, q_ref, ok) of the curve solver's hand-scheduled division on the device (include/omo.h)."""
        num = np.ascontiguousarray(num, dtype=np.float64)
        den = np.ascontiguousarray(den, dtype=np.float64)
        n = num.size
        qf, qr, ok = np.empty(n), np.empty(n), np.empty(n, dtype=np.int32)
        self._ck(_omo.omo_selftest_division(self.h, num.ctypes.data_as(_dp), den.ctypes.data_as(_dp), n, qf.ctypes.data_as(_dp),
                                            qr.ctypes.data_as(_dp), ok.ctypes.data_as(_ip)), "omo_selftest_division")
        return qf, qr, ok

    def grad_sparsity(self):
        """(work items executed, work items of the dense evaluation) of the last finite-difference call."""
        out = (_ull * 2)()
        self._ck(_omo.omo_grad_sparsity(self.h, out), "omo_grad_sparsity")
        return int(out[0]), int(out[1])

    def dfma_peak(self):
        v = C.c_double(0.0)
        self._ck(_omo.omo_measure_dfma_peak(self.h, C.byref(v)), "omo_measure_dfma_peak")
        return v.value


class PeerGather:
    """omo_peer_gather (include/omo.h): the f values of ONE batch evaluated by several PROCESSES, one GPU each, gathered into every
    rank's device memory by the kernel that forms them (NVLink / NVSwitch peer stores, no NCCL call). `exchange(handle) -> [handles in
    rank order]` is the caller's transport (torch.distributed.all_gather_object, MPI, files)."""
    HANDLE_BYTES = 64

    def __init__(self, engine, rank, world, capacity, exchange=None):
        """Creates this rank's gathered array. With `exchange` the handles are swapped and the peers mapped right away; without it the
        caller passes `self.handle` through its own collective and calls connect(handles) -- the form to use when a rank may fail to
        create its array: every rank then still takes part in the collective (bench.py does this)."""
        self.eng, self.rank, self.world, self.capacity = engine, rank, world, capacity
        h = _vp()
        buf = C.create_string_buffer(self.HANDLE_BYTES)
        engine._ck(_omo.omo_peer_gather_create(engine.h, rank, world, capacity, C.byref(h), buf), "omo_peer_gather_create")
        self.h = h.value
        self.handle = buf.raw
        if exchange is not None:
            self.connect(exchange(self.handle))

    def connect(self, handles):
        if len(handles) != self.world or any(x is None or len(x) != self.HANDLE_BYTES for x in handles):
            raise EngineError("peer gather: need one 64-byte handle per rank")
        self.eng._ck(_omo.omo_peer_gather_connect(self.h, b"".join(handles)), "omo_peer_gather_connect")

    def eval(self, dX, B_local, first, sync=True):
        """Evaluates this rank's slice [first, first + B_local) of the batch; afterwards values() holds every rank's f."""
        self.eng._ck(_omo.omo_peer_gather_eval(self.h, dX, B_local, first, int(sync)), "omo_peer_gather_eval")

    def values(self):
        """Device address of the gathered f (capacity doubles) of the last call."""
        return _omo.omo_peer_gather_values(self.h)

    def error(self):
        return _omo.omo_peer_gather_error(self.h)

    def close(self):
        if self.h:
            _omo.omo_peer_gather_destroy(self.h)
            self.h = None


class MultiEngine:
    """Raw view of an omo_multi* (include/omo.h) owned by a Core: one host thread, several GPUs of one box."""

    GATHER_HOST, GATHER_NCCL = 0, 1

    def __init__(self, handle):
        if not handle:
            raise EngineError("multi-device engine is not available: " + (_core.optimo_last_error() or b"").decode())
        self.h = handle

    def _ck(self, st, what):
        if st != 0:
            raise EngineError(f"{what}: status {st}: {(_omo.omo_multi_last_error(self.h) or b'').decode()}")

    @property
    def num_devices(self):
        return _omo.omo_multi_num_devices(self.h)

    def devices(self):
        return [_omo.omo_multi_device(self.h, i) for i in range(self.num_devices)]

    def ctx(self, i):
        return Engine(_omo.omo_multi_ctx(self.h, i))

    def set_gather(self, mode):
        self._ck(_omo.omo_multi_set_gather(self.h, int(mode)), "omo_multi_set_gather")

    @property
    def gather(self):
        return _omo.omo_multi_get_gather(self.h)

    def eval_host_ptr(self, x_ptr, B, f_ptr):
        self._ck(_omo.omo_multi_eval(self.h, C.cast(x_ptr, _dp), B, C.cast(f_ptr, _dp), None), "omo_multi_eval")

    def eval(self, X, want_terms=False):
        X, xp = _d(X)
        D = self.ctx(0).D
        X2 = X.reshape(-1, D)
        f = np.zeros(X2.shape[0])
        terms = np.zeros((X2.shape[0], 3)) if want_terms else None
        self._ck(_omo.omo_multi_eval(self.h, xp, X2.shape[0], f.ctypes.data_as(_dp), terms.ctypes.data_as(_dp) if want_terms else None), "omo_multi_eval")
        return (f, terms) if want_terms else f

    def eval_free_host_ptr(self, x_base, free_idx, xf_ptr, B, f_ptr):
        xb, xbp = _d(x_base)
        idx, ip = _i(free_idx)
        self._ck(_omo.omo_multi_eval_free(self.h, xbp, ip, idx.size, C.cast(xf_ptr, _dp), B, C.cast(f_ptr, _dp), None), "omo_multi_eval_free")

    def eval_grad(self, x, free_idx, h=1e-4):
        x, xp = _d(x)
        idx, ip = _i(free_idx)
        g, fpm = np.zeros(idx.size), np.zeros((idx.size, 2))
        f0 = C.c_double(0.0)
        self._ck(_omo.omo_multi_eval_grad(self.h, xp, ip, idx.size, float(h), C.byref(f0), g.ctypes.data_as(_dp), fpm.ctypes.data_as(_dp)), "omo_multi_eval_grad")
        return f0.value, g, fpm


class Core:
    """Headless Core (one instance == one scene + one lazily created GPU engine)."""

    def __init__(self, scene=None, device=None, devices=None):
        self.h = _core.optimo_core_new()
        if not self.h:
            raise EngineError("optimo_core_new failed")
        if device is not None:
            self._ck(_core.optimo_set_device(self.h, int(device)))
        if devices is not None:
            self.SetDevices(devices)
        if scene is not None:
            if isinstance(scene, dict):
                self.ReadJsonString(json.dumps(scene))
            else:
                self.ReadJson(scene)

    def __del__(self):
        try:
            if self.h:
                _core.optimo_core_free(self.h)
                self.h = None
        except Exception:
            pass

    def _ck(self, rc):
        if rc == 0:
            return
        msg = (_core.optimo_last_error() or b"").decode()
        if rc == 2:
            raise ForcedStop(msg)
        raise EngineError(msg)

    def SetDevices(self, devices):
        """Several GPUs of one box behind the same calls (batches sharded over candidates, gradients over columns)."""
        a, p = _i(list(devices))
        self._ck(_core.optimo_set_devices(self.h, p, a.size))

    def multi_engine(self):
        return MultiEngine(_core.optimo_engine_multi(self.h))

    # ---- IO / structure ------------------------------------------------------------------------------
    def ReadJson(self, path):
        self._ck(_core.optimo_read_json(self.h, os.fspath(path).encode()))

    def ReadJsonString(self, text):
        self._ck(_core.optimo_read_json_string(self.h, text.encode()))

    def WriteJson(self, path):
        self._ck(_core.optimo_write_json(self.h, os.fspath(path).encode()))

    def scene_json(self):
        n = _core.optimo_scene_json(self.h, None, 0)
        buf = C.create_string_buffer(n + 1)
        _core.optimo_scene_json(self.h, buf, n + 1)
        return buf.value.decode()

    @property
    def min_frame_(self):
        return _core.optimo_min_frame(self.h)

    @property
    def max_frame_(self):
        return _core.optimo_max_frame(self.h)

    @property
    def window(self):
        """[min_frame - 2, max_frame + 2], the objective window of PerformOptimization (core_opt.cpp:125-126)."""
        return self.min_frame_ - 2.0, self.max_frame_ + 2.0

    @property
    def J(self):
        return _core.optimo_num_joints(self.h)

    @property
    def D(self):
        return _core.optimo_descriptor_size(self.h)

    def item_names(self):
        out = []
        buf = C.create_string_buffer(256)
        for k in range(_core.optimo_num_items(self.h)):
            _core.optimo_item_name(self.h, k, buf, 256)
            out.append(buf.value.decode())
        return out

    def GetDescriptor(self):
        x = np.zeros(self.D)
        self._ck(_core.optimo_get_descriptor(self.h, x.ctypes.data_as(_dp)))
        return x

    def SetDescriptor(self, x):
        a, p = _d(x)
        self._ck(_core.optimo_set_descriptor(self.h, p, a.size))

    def set_cyclic(self, flag):
        _core.optimo_set_cyclic(self.h, int(bool(flag)))

    def joint_physical(self, joint):
        out = np.zeros(13)
        self._ck(_core.optimo_joint_physical(self.h, joint, out.ctypes.data_as(_dp)))
        return out[0], out[1:4].copy(), out[4:13].reshape(3, 3).copy()

    # ---- runtime state -------------------------------------------------------------------------------
    def UpdateOriginalParameters(self, x):
        a, p = _d(x)
        self._ck(_core.optimo_update_original_parameters(self.h, p, a.size))

    def RetrieveOriginalParameters(self):
        x = np.zeros(self.D)
        self._ck(_core.optimo_retrieve_original_parameters(self.h, x.ctypes.data_as(_dp)))
        return x

    def UpdateGlobalWeight(self, w):
        _core.optimo_update_global_weight(self.h, float(w))

    def UpdatePerItemWeight(self, w, item_name):
        self._ck(_core.optimo_update_per_item_weight(self.h, float(w), item_name.encode()))

    def set_kernels(self, kernels):
        _core.optimo_clear_kernels(self.h)
        for amp, mu, sigma in kernels:
            _core.optimo_add_kernel(self.h, float(amp), float(mu), float(sigma))

    def GetCostWeight(self, t):
        return _core.optimo_get_cost_weight(self.h, float(t))

    def CalculateParameterWeights(self):
        w = np.zeros(self.D)
        self._ck(_core.optimo_parameter_weights(self.h, w.ctypes.data_as(_dp)))
        return w

    def select_items(self, names):
        for n in self.item_names():
            self._ck(_core.optimo_set_selected(self.h, n.encode(), int(n in names)))
        _core.optimo_update_freezed_dofs_from_selection(self.h)

    @property
    def freezed_dofs_(self):
        f = np.zeros(self.D, dtype=np.int32)
        _core.optimo_get_freezed_dofs(self.h, f.ctypes.data_as(_ip))
        return f

    @freezed_dofs_.setter
    def freezed_dofs_(self, f):
        a, p = _i(f)
        self._ck(_core.optimo_set_freezed_dofs(self.h, p, a.size))

    def free_indices(self):
        return np.nonzero(self.freezed_dofs_ == 0)[0].astype(np.int32)

    # ---- the hot path --------------------------------------------------------------------------------
    def CalculateFullObjective(self, min_t, max_t, x, x_original, want_terms=False):
        xa, xp = _d(x)
        oa, op = _d(x_original)
        f = C.c_double(0.0)
        terms = np.zeros(3)
        self._ck(_core.optimo_calculate_full_objective(self.h, min_t, max_t, xp, op, xa.size, C.byref(f), terms.ctypes.data_as(_dp)))
        return (f.value, terms) if want_terms else f.value

    def CalculateFullObjectiveBatch(self, min_t, max_t, X, x_original, want_terms=False):
        oa, op = _d(x_original)
        Xa, Xp = _d(X)
        Xa = Xa.reshape(-1, oa.size)
        f = np.zeros(Xa.shape[0])
        terms = np.zeros((Xa.shape[0], 3)) if want_terms else None
        self._ck(_core.optimo_calculate_full_objective_batch(self.h, min_t, max_t, Xp, Xa.shape[0], op, oa.size, f.ctypes.data_as(_dp),
                                                            terms.ctypes.data_as(_dp) if want_terms else None))
        return (f, terms) if want_terms else f

    def CalculateGradient(self, min_t, max_t, x, x_original, free_idx=None, h=1e-4):
        """Returns (f0, grad[F], fpm[F,2])."""
        xa, xp = _d(x)
        oa, op = _d(x_original)
        idx, ip = _i(self.free_indices() if free_idx is None else free_idx)
        g, fpm = np.zeros(idx.size), np.zeros((idx.size, 2))
        f0 = C.c_double(0.0)
        self._ck(_core.optimo_calculate_gradient(self.h, min_t, max_t, xp, op, xa.size, ip, idx.size, float(h), C.byref(f0),
                                                g.ctypes.data_as(_dp), fpm.ctypes.data_as(_dp)))
        return f0.value, g, fpm

    def CalculateAnalyticGradient(self, min_t, max_t, x, x_original, free_idx=None):
        """Next row N4: (f0, grad[F], analytic_used) -- the analytic gradient, or the finite-difference one where it does not apply."""
        xa, xp = _d(x)
        oa, op = _d(x_original)
        idx, ip = _i(self.free_indices() if free_idx is None else free_idx)
        g = np.zeros(idx.size)
        f0, used = C.c_double(0.0), C.c_int(0)
        self._ck(_core.optimo_calculate_analytic_gradient(self.h, min_t, max_t, xp, op, xa.size, ip, idx.size, C.byref(f0), g.ctypes.data_as(_dp), C.byref(used)))
        return f0.value, g, bool(used.value)

    def debug_scene_table(self, nbytes):
        """Host-only: (bytes of the flattened device scene table, sub_end[J]) for the CPU check of the analytic gradient's arithmetic."""
        buf = C.create_string_buffer(nbytes)
        sub_end = np.zeros(self.J, dtype=np.int32)
        self._ck(_core.optimo_debug_scene_table(self.h, C.cast(buf, _vp), nbytes, sub_end.ctypes.data_as(_ip)))
        return buf, sub_end

    def positions(self, min_t, max_t, x):
        xa, xp = _d(x)
        n = _core.optimo_num_samples(self.h, min_t, max_t)
        p = np.zeros((n, 3 * self.J))
        self._ck(_core.optimo_positions(self.h, min_t, max_t, xp, xa.size, p.ctypes.data_as(_dp)))
        return p

    # ---- next rows N2 / N3 ----------------------------------------------------------------------
    def CalculateTorques(self, t, want_states=False):
        """Core::CalculateTorques (core_rbdl.cpp:264-289) of the current object. A scalar t gives tau[dof]; a sequence of
        M times is ONE launch and gives tau[M, dof]. want_states: also (q, qd, qdd) of SetCurrentState (:168-198)."""
        scalar = np.isscalar(t)
        ta, tp = _d(np.atleast_1d(t))
        dof = 3 * self.J
        outs = [np.zeros((ta.size, dof)) for _ in range(4 if want_states else 1)]
        ptrs = [o.ctypes.data_as(_dp) for o in outs] + [None] * (4 - len(outs))
        self._ck(_core.optimo_calculate_torques(self.h, tp, ta.size, *ptrs))
        if scalar:
            outs = [o[0] for o in outs]
        return tuple(outs) if want_states else outs[0]

    def CalculateEulerAngles(self, times):
        """Joint::GetEulerAngle(t) (joint.cpp:36-52) of every joint, root included: [M, 3J]."""
        ta, tp = _d(np.atleast_1d(times))
        e = np.zeros((ta.size, 3 * self.J))
        self._ck(_core.optimo_euler_angles(self.h, tp, ta.size, e.ctypes.data_as(_dp)))
        return e

    def UpdateMaximumTorqueAndCost(self):
        a, b = C.c_double(0.0), C.c_double(0.0)
        self._ck(_core.optimo_update_maximum_torque_and_cost(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def ExportMotion(self, path):
        self._ck(_core.optimo_export_motion(self.h, str(path).encode()))

    def motion_json(self):
        n = _core.optimo_motion_json(self.h, None, 0)
        if n < 0:
            self._ck(1)
        buf = C.create_string_buffer(n + 1)
        _core.optimo_motion_json(self.h, buf, n + 1)
        return buf.value.decode()

    def PushHistory(self):
        self._ck(_core.optimo_push_history(self.h))

    def PopHistory(self):
        self._ck(_core.optimo_pop_history(self.h))

    def ClearHistory(self):
        _core.optimo_clear_history(self.h)

    def GetHistorySize(self):
        return _core.optimo_history_size(self.h)

    def WriteHistory(self, path):
        self._ck(_core.optimo_write_history(self.h, str(path).encode()))

    def EvaluateObjective(self, x_free, x_initial, min_t=None, max_t=None, want_gradient=False, h=1e-4):
        lo, hi = self.window
        min_t = lo if min_t is None else min_t
        max_t = hi if max_t is None else max_t
        xa, xp = _d(x_free)
        ia, ip = _d(x_initial)
        f = C.c_double(0.0)
        g = np.zeros(xa.size) if want_gradient else None
        self._ck(_core.optimo_evaluate_objective(self.h, xp, xa.size, ip, ia.size, min_t, max_t, int(want_gradient), float(h), C.byref(f),
                                                g.ctypes.data_as(_dp) if want_gradient else None))
        return (f.value, g) if want_gradient else f.value

    @property
    def stop_optimization_(self):
        return bool(_core.optimo_get_stop_optimization(self.h))

    @stop_optimization_.setter
    def stop_optimization_(self, flag):
        _core.optimo_set_stop_optimization(self.h, int(bool(flag)))

    def PerformOptimization(self, explicit_termination=False):
        """Returns (final objective value, function evaluations)."""
        v, n = C.c_double(0.0), C.c_int(0)
        self._ck(_core.optimo_perform_optimization(self.h, int(explicit_termination), C.byref(v), C.byref(n)))
        return v.value, n.value

    def set_batched_start(self, on):
        """Batched BOBYQA start: the solver's first 2M+1 points as one GPU batch (default on)."""
        _core.optimo_set_batched_start(self.h, int(bool(on)))

    @property
    def last_solve_batched(self):
        return _core.optimo_last_solve_batched(self.h)

    def PerformOptimizationWithGradient(self, explicit_termination=False, algorithm="LD_LBFGS", analytic=False):
        """Gradient-based solve (finite-difference gradient on the GPU, or the analytic one). Returns (final objective value, function evaluations)."""
        v, n = C.c_double(0.0), C.c_int(0)
        self._ck(_core.optimo_perform_optimization_gradient(self.h, int(explicit_termination), (algorithm + ("+analytic" if analytic else "")).encode(), C.byref(v), C.byref(n)))
        return v.value, n.value

    def probe_specialised_kernel(self, fk=False, count=False, dump_prefix=None, self_preparing=False):
        """Host-only: generates and NVRTC-compiles the scene-specialised kernel (self_preparing: the one-launch instance for
        single candidates and small gradients); returns (ok, log)."""
        buf = C.create_string_buffer(1 << 16)
        rc = _core.optimo_probe_specialised_kernel(self.h, int(fk), int(bool(count)) | (2 if self_preparing else 0), dump_prefix.encode() if dump_prefix else None, buf, len(buf))
        return rc == 0, buf.value.decode(errors="replace")

    def record_solve_trace(self, on=True):
        _core.optimo_record_solve_trace(self.h, int(bool(on)))

    def solve_trace(self):
        """(X_free [rows, M], f [rows]) of every callback of the last recorded solve, in call order."""
        rows, dim = C.c_int(0), C.c_int(0)
        self._ck(_core.optimo_solve_trace_size(self.h, C.byref(rows), C.byref(dim)))
        a = np.zeros((rows.value, dim.value + 1))
        if a.size:
            self._ck(_core.optimo_get_solve_trace(self.h, a.ctypes.data_as(_dp)))
        return a[:, :-1].copy(), a[:, -1].copy()

    def GetOptimizationCache(self):
        n, dof = C.c_int(0), C.c_int(0)
        self._ck(_core.optimo_cache_size(self.h, C.byref(n), C.byref(dof)))
        out = {}
        for which, (name, width) in enumerate([("time", 0), ("q", dof.value), ("qd", dof.value), ("qdd", dof.value), ("tau", dof.value), ("local_cost", 0)]):
            a = np.zeros((n.value, width)) if width else np.zeros(n.value)
            self._ck(_core.optimo_get_cache(self.h, which, a.ctypes.data_as(_dp)))
            out[name] = a
        return out

    def ClearOptimizationCache(self):
        _core.optimo_clear_optimization_cache(self.h)

    def engine(self, synced_window=None):
        """The omo_ctx behind this Core. With `synced_window=(min_t, max_t)` the window / weights / x_original
        are pushed first (by one evaluation of the current descriptor)."""
        if synced_window is None:
            h = _core.optimo_engine(self.h)
        else:
            h = _core.optimo_engine_synced(self.h, float(synced_window[0]), float(synced_window[1]))
        return Engine(h)

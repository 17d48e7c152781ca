"""TEST INFRASTRUCTURE: drives the reference's own NLopt (third_party/_ref/libnlopt.so, compiled from
/root/reference/external/nlopt) from Python with an arbitrary objective, replaying what Core::PerformOptimization
(OptiMo/core_opt.cpp:60-157) + nloptutil::solve (external/nlopt-util/include/nlopt-util.hpp:44-195) set up:
LN_BOBYQA, box bounds, maxeval / ftol_rel / xtol_rel, NLopt's default initial step scaled by 0.06."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "third_party", "_ref", "libnlopt.so")
_FUNC = C.CFUNCTYPE(C.c_double, C.c_uint, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)


def available():
    return os.path.exists(LIB)


def bounds_like_perform_optimization(x_initial):
    """core_opt.cpp:82-97: x +- 6, theta in +-0.4 pi, time_backward <= -1e-6, time_forward >= +1e-6; clamped start."""
    x = np.asarray(x_initial, dtype=np.float64).copy()
    upper, lower = x + 6.0, x - 6.0
    upper[2::5] = +np.pi * 0.4
    lower[2::5] = -np.pi * 0.4
    upper[3::5] = -1e-06
    lower[4::5] = +1e-06
    return np.maximum(np.minimum(x, upper), lower), lower, upper


def solve_bobyqa(fun, x_start, lower, upper, maxeval=2000, ftol_rel=1e-7, xtol_rel=1e-7, step_scale=0.06):
    """Returns (x_star, f_star, n_evals, [(x, f), ...] trace)."""
    lib = C.CDLL(LIB)
    lib.nlopt_create.restype = C.c_void_p
    lib.nlopt_create.argtypes = [C.c_int, C.c_uint]
    lib.nlopt_algorithm_from_string.argtypes = [C.c_char_p]
    dp = C.POINTER(C.c_double)
    for name, args in [("nlopt_set_lower_bounds", [C.c_void_p, dp]), ("nlopt_set_upper_bounds", [C.c_void_p, dp]),
                       ("nlopt_set_maxeval", [C.c_void_p, C.c_int]), ("nlopt_set_ftol_rel", [C.c_void_p, C.c_double]),
                       ("nlopt_set_xtol_rel", [C.c_void_p, C.c_double]), ("nlopt_get_initial_step", [C.c_void_p, dp, dp]),
                       ("nlopt_set_initial_step", [C.c_void_p, dp]), ("nlopt_set_min_objective", [C.c_void_p, _FUNC, C.c_void_p]),
                       ("nlopt_optimize", [C.c_void_p, dp, dp]), ("nlopt_get_numevals", [C.c_void_p]), ("nlopt_destroy", [C.c_void_p])]:
        getattr(lib, name).argtypes = args
    n = len(x_start)
    trace = []

    def cb(nn, xp, gp, _):
        x = np.ctypeslib.as_array(xp, shape=(nn,)).copy()
        f = float(fun(x))
        trace.append((x, f))
        return f

    cfun = _FUNC(cb)
    opt = lib.nlopt_create(lib.nlopt_algorithm_from_string(b"LN_BOBYQA"), n)
    as_p = lambda a: a.ctypes.data_as(dp)
    lo, up = np.ascontiguousarray(lower, dtype=np.float64), np.ascontiguousarray(upper, dtype=np.float64)
    lib.nlopt_set_upper_bounds(opt, as_p(up))
    lib.nlopt_set_lower_bounds(opt, as_p(lo))
    lib.nlopt_set_maxeval(opt, maxeval)
    lib.nlopt_set_ftol_rel(opt, ftol_rel)
    lib.nlopt_set_xtol_rel(opt, xtol_rel)
    lib.nlopt_set_min_objective(opt, cfun, None)
    x = np.ascontiguousarray(x_start, dtype=np.float64).copy()
    step = np.zeros(n)
    lib.nlopt_get_initial_step(opt, as_p(x), as_p(step))
    step *= step_scale
    lib.nlopt_set_initial_step(opt, as_p(step))
    fval = C.c_double(0.0)
    lib.nlopt_optimize(opt, as_p(x), C.byref(fval))
    evals = lib.nlopt_get_numevals(opt)
    lib.nlopt_destroy(opt)
    return x, fval.value, evals, trace

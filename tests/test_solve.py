"""BASELINE.json configs[1]: resources/scene-fox-tail.json, full NLopt (LN_BOBYQA) solve with the GPU objective
callback -- Core::PerformOptimization (OptiMo/core_opt.cpp:60-157) over nloptutil::solve.

Parity is by REPLAY: every x the solver asked the GPU callback about is re-evaluated
by the CPU oracle and f must agree to 1e-9 relative; the end value is compared loosely with an oracle-driven solve
through the same NLopt build (BOBYQA trajectories fork on last-bit differences and both stop at the maxeval cap)."""
import json
import os

import numpy as np
import pytest

import nlopt_driver
from conftest import GOLDEN, scene_path

SOLVE_GOLDEN = os.path.join(GOLDEN, "solve_fox-tail.json")
TAIL = ["tail_root", "tail_center"]
needs_nlopt = pytest.mark.skipif(not nlopt_driver.available(), reason="third_party/_ref/libnlopt.so not built (make -C third_party)")


def _oracle_solve(O):
    orc = O.Oracle(scene_path("fox-tail"), "port")
    orc.select_items(TAIL)
    idx = orc.free_indices()
    x_init, lower, upper = nlopt_driver.bounds_like_perform_optimization(orc.descriptor())
    x_orig = orc.original()

    def fun(xf):
        full = x_init.copy()
        full[idx] = xf
        return orc.full_objective(full, x_orig)[0]

    xs, fs, evals, trace = nlopt_driver.solve_bobyqa(fun, x_init[idx], lower[idx], upper[idx])
    return orc, idx, x_init, xs, fs, evals, trace


@needs_nlopt
def test_oracle_driven_solve_matches_committed_golden(oracle_port):
    """Pins the CPU side of the replay: same NLopt sources + same oracle => same trajectory as when the golden was made."""
    orc, idx, x_init, xs, fs, evals, trace = _oracle_solve(oracle_port)
    assert idx.size == 27
    assert fs < trace[0][1] * 0.9          # it really minimised something
    assert fs == min(f for _, f in trace)
    with open(SOLVE_GOLDEN) as fh:
        g = json.load(fh)
    assert abs(trace[0][1] - g["f_initial"]) <= 1e-12 * abs(g["f_initial"])
    assert abs(fs - g["f_final"]) <= 1e-6 * abs(g["f_final"])
    assert abs(evals - g["evals"]) <= max(20, g["evals"] // 5)


@pytest.mark.gpu
@needs_nlopt
def test_fox_tail_solve_replay_parity(oracle_port):
    from optimo_b200.core import Core

    c = Core(scene_path("fox-tail"))
    c.select_items(TAIL)
    x_before = c.GetDescriptor()
    c.record_solve_trace(True)
    f_final, evals = c.PerformOptimization(False)
    X, F = c.solve_trace()
    assert X.shape[1] == 27 and X.shape[0] == evals and evals > 50

    orc = oracle_port.Oracle(scene_path("fox-tail"), "port")
    orc.select_items(TAIL)
    idx = orc.free_indices()
    x_init, _, _ = nlopt_driver.bounds_like_perform_optimization(x_before)
    x_orig = orc.original()
    worst = 0.0
    for xf, f in zip(X, F):
        full = x_init.copy()
        full[idx] = xf
        fo = orc.full_objective(full, x_orig)[0]
        worst = max(worst, abs(f - fo) / abs(fo))
    assert worst <= 1e-9, worst

    assert f_final == F.min()
    with open(SOLVE_GOLDEN) as fh:
        g = json.load(fh)
    assert abs(F[0] - g["f_initial"]) <= 1e-9 * abs(g["f_initial"])
    # End value: both solves stop at the maxeval = 2000 cap (core_opt.cpp:132) still descending, and BOBYQA's path forks
    # on last-bit differences of f (measured from 26566.8: 2693.38 oracle-driven; 2669.40 and 4302.75 GPU-driven with two
    # kernel versions whose f agree with the oracle to 1e-13), so only a substantial descent is asserted here; the
    # per-evaluation replay above is the parity statement.
    assert f_final < 0.5 * F[0]
    # the solve leaves x_opt in the scene (core_opt.cpp:156)
    x_after = c.GetDescriptor()
    assert np.array_equal(x_after[idx], X[np.argmin(F)])
    frozen = np.setdiff1d(np.arange(x_after.size), idx)
    assert np.array_equal(x_after[frozen], x_init[frozen])


@pytest.mark.gpu
@needs_nlopt
@pytest.mark.parametrize("algorithm", ["LD_LBFGS", "LD_SLSQP"])
def test_fox_tail_gradient_based_solve(oracle_port, algorithm):
    """Next row N1: the fox-tail solve driven by a gradient-based NLopt algorithm; every callback's f and
    finite-difference gradient come from one sparse GPU batch (omo_eval_grad). The values the solver saw are replayed on the
    CPU oracle; the descent must at least match what 2000 BOBYQA evaluations reach (f_final of the committed golden)."""
    from optimo_b200.core import Core

    c = Core(scene_path("fox-tail"))
    c.select_items(TAIL)
    x_before = c.GetDescriptor()
    c.record_solve_trace(True)
    f_final, evals = c.PerformOptimizationWithGradient(False, algorithm)
    X, F = c.solve_trace()
    assert X.shape[1] == 27 and 5 < evals <= 2000

    orc = oracle_port.Oracle(scene_path("fox-tail"), "port")
    orc.select_items(TAIL)
    idx = orc.free_indices()
    x_orig = orc.original()
    x_init = x_before.copy()
    worst = 0.0
    for xf, f in list(zip(X, F))[:: max(1, len(F) // 40)]:
        full = x_init.copy()
        full[idx] = xf
        fo = orc.full_objective(full, x_orig)[0]
        worst = max(worst, abs(f - fo) / abs(fo))
    assert worst <= 1e-9, worst
    with open(SOLVE_GOLDEN) as fh:
        g = json.load(fh)
    assert f_final <= F[0] and f_final == F.min()
    assert f_final < 1.5 * g["f_final"], (f_final, g["f_final"], evals)
    # handles stayed on their side of zero with room for the difference step
    x_after = c.GetDescriptor().reshape(-1, 5)
    assert (x_after[:, 3] < 0).all() and (x_after[:, 4] > 0).all()
    with pytest.raises(Exception):
        c.PerformOptimizationWithGradient(False, "LN_BOBYQA")


@needs_nlopt
@pytest.mark.gpu
def test_fox_tail_solve_over_the_analytic_gradient(oracle_port):
    """Next row N4 in the solver: LD_SLSQP driven by omo_eval_grad_analytic reaches the optimum of the finite-difference-driven
    solve (same bounds, same tolerances) -- and every callback's f is the reference objective."""
    from optimo_b200.core import Core

    def run(analytic):
        c = Core(scene_path("fox-tail"))
        c.select_items(["tail_root", "tail_center"])
        c.record_solve_trace(True)
        f, evals = c.PerformOptimizationWithGradient(False, "LD_SLSQP", analytic=analytic)
        return c, f, evals
    c_fd, f_fd, n_fd = run(False)
    c_an, f_an, n_an = run(True)
    assert abs(f_an - f_fd) <= 1e-4 * abs(f_fd), (f_an, f_fd)
    assert n_an <= 2 * n_fd
    X, F = c_an.solve_trace()
    o = oracle_port.Oracle(scene_path("fox-tail"), "port"); o.set_sampling_concurrency(1)
    o.select_items(["tail_root", "tail_center"])
    x0 = o.descriptor()
    idx = o.free_indices()
    x_init, _, _ = nlopt_driver.bounds_like_perform_optimization(x0)
    for k in range(0, X.shape[0], 7):
        full = x_init.copy(); full[idx] = X[k]
        fo, _ = o.full_objective(full, x0)
        assert abs(F[k] - fo) <= 1e-9 * abs(fo)


@needs_nlopt
@pytest.mark.gpu
@pytest.mark.parametrize("name,items", [("fox-tail", ["tail_root", "tail_center"]), ("stick-figure", None)])
def test_batched_bobyqa_start_leaves_the_iterates_unchanged(name, items):
    """SURVEY.md section 8(f) N1, second half: BOBYQA's first 2M+1 interpolation points (bobyqa.c prelim_: x0 and x0 +- rho e_i, shifted at
    bounds, in NLopt's rescaled variables) are predicted bit for bit, evaluated as one batch and served from a cache. Every one of
    them must be hit, and the solve must ask the same questions and get the same answers as the serial start."""
    from optimo_b200.core import Core

    def run(batched):
        c = Core(scene_path(name))
        if items:
            c.select_items(items)
        c.set_batched_start(batched)
        c.record_solve_trace(True)
        f, evals = c.PerformOptimization(False)
        X, F = c.solve_trace()
        return f, evals, X, F, c.last_solve_batched, c.free_indices().size

    f0, e0, X0, F0, h0, M = run(False)
    f1, e1, X1, F1, h1, _ = run(True)
    assert h0 == 0 and h1 == 2 * M + 1, (h1, M)
    assert (f0, e0) == (f1, e1)
    assert np.array_equal(X0, X1) and np.array_equal(F0, F1)

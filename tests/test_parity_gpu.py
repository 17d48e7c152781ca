"""Parity of the CUDA path (through the C ABI / headless Core) against the CPU oracle and the committed goldens.

Bar (BASELINE.json north_star): objective and gradient values within 1e-9 relative in FP64. The per-stage
arrays (q, qd, qdd, tau, local_cost, positions) are compared much tighter -- they only differ by FMA
contraction and libm-vs-libdevice last-ulp effects.
"""
import numpy as np
import pytest

from conftest import SCENES, all_cases, configure, elem_err, gradient_bound, load_cases, load_scene, rel_err, scene_path

from optimo_b200 import synthetic
from optimo_b200.core import Core, EngineError, ForcedStop, grad_from_columns, shard_range

pytestmark = pytest.mark.gpu

TOL = 1e-9       # the contract
TIGHT = 2e-11    # what the kernels actually achieve on smooth stages


def _product_for(name, case):
    c = Core(scene_path(name))
    configure(c, case, False)
    return c


@pytest.mark.parametrize("name,tag", all_cases())
def test_objective_terms_and_cache_match_goldens(name, tag):
    cases, g = load_cases(name)
    case = next(k for k in cases if k["tag"] == tag)
    c = _product_for(name, case)
    x, x0 = g[f"{tag}/x"], g[f"{tag}/x_original"]
    c.UpdateOriginalParameters(x0)
    lo, hi = c.window
    f, terms = c.CalculateFullObjective(lo, hi, x, x0, want_terms=True)
    fg = float(g[f"{tag}/f"])
    assert abs(f - fg) <= TOL * abs(fg), (f, fg)
    assert np.allclose(terms, g[f"{tag}/terms"], rtol=TOL, atol=1e-12)
    cache = c.GetOptimizationCache()
    assert np.array_equal(cache["time"], g[f"{tag}/time"])          # the accumulated t += 0.2 grid, bit for bit
    for k in ("q", "qd", "qdd", "tau", "local_cost"):
        assert cache[k].shape == g[f"{tag}/{k}"].shape, k
        tol = TIGHT if "ik" not in tag else 1e-9                       # IK angles are themselves converged to 1e-6 only
        assert rel_err(cache[k], g[f"{tag}/{k}"]) <= tol, (k, rel_err(cache[k], g[f"{tag}/{k}"]))
        assert elem_err(cache[k], g[f"{tag}/{k}"], tol) <= 1.0, (k, elem_err(cache[k], g[f"{tag}/{k}"], tol))   # entry by entry
    assert rel_err(c.positions(lo, hi, x), g[f"{tag}/positions"]) <= (TIGHT if "ik" not in tag else 1e-9)
    assert elem_err(c.positions(lo, hi, x), g[f"{tag}/positions"], TIGHT if "ik" not in tag else 1e-9) <= 1.0
    # the reference leaves the object at x (core_rbdl.cpp:371)
    xx = c.GetDescriptor()
    if not case["cyclic"]:
        assert np.array_equal(xx, x)


@pytest.mark.parametrize("name,tag", all_cases())
def test_finite_difference_gradient_matches_goldens(name, tag):
    cases, g = load_cases(name)
    case = next(k for k in cases if k["tag"] == tag)
    c = _product_for(name, case)
    x, x0 = g[f"{tag}/x"], g[f"{tag}/x_original"]
    c.UpdateOriginalParameters(x0)
    lo, hi = c.window
    assert np.array_equal(c.free_indices(), g[f"{tag}/free_idx"])
    f0, grad, fpm = c.CalculateGradient(lo, hi, x, x0, h=1e-4)
    assert abs(f0 - float(g[f"{tag}/f0"])) <= TOL * abs(float(g[f"{tag}/f0"]))
    assert np.max(np.abs(fpm - g[f"{tag}/fpm"]) / np.abs(g[f"{tag}/fpm"])) <= TOL       # every perturbed value
    gg = g[f"{tag}/grad"]
    assert np.max(np.abs(grad - gg)) <= gradient_bound(gg, f0, 1e-4), (np.max(np.abs(grad - gg)), gradient_bound(gg, f0, 1e-4))


@pytest.mark.parametrize("name", SCENES)
def test_batch_matches_oracle_on_seeded_candidates(oracle_port, name):
    scene = load_scene(name)
    o = oracle_port.Oracle(scene, "port")
    o.set_sampling_concurrency(1)
    c = Core(scene_path(name))
    o.set_weight_trajectory(5.0); c.UpdateGlobalWeight(5.0)
    x0 = c.GetDescriptor()
    B = 24
    X = synthetic.perturbed_candidates(x0, B, seed_base=0xB200)
    lo, hi = c.window
    f, terms = c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)
    for b in range(B):
        fo, to = o.full_objective(X[b], x0)
        assert abs(f[b] - fo) <= TOL * abs(fo), (b, f[b], fo)
        assert np.allclose(terms[b], to, rtol=TOL, atol=1e-12)
    # single-candidate API == row of the batch: the one candidate goes through the self-preparing scene-specialised kernel, the
    # 24-row batch through the table-driven one (small launch); the two agree to rounding of the structural zeros
    assert abs(c.CalculateFullObjective(lo, hi, X[5], x0) - f[5]) <= 1e-13 * abs(f[5])


def test_evaluate_objective_callback_and_cancel():
    """EvaluateObjective (core_opt.cpp:42-57): free -> full expansion, forced_stop on the cancel flag."""
    c = Core(scene_path("stick-figure"))
    x0 = c.GetDescriptor()
    idx = c.free_indices()
    assert idx.size == 78
    rng = np.random.default_rng(5)
    xf = x0[idx] + 0.01 * rng.standard_normal(idx.size)
    full = x0.copy(); full[idx] = xf
    lo, hi = c.window
    f = c.EvaluateObjective(xf, x0)
    assert f == c.CalculateFullObjective(lo, hi, full, x0)
    f2, g = c.EvaluateObjective(xf, x0, want_gradient=True)
    assert f2 == f
    _, g_ref, _ = c.CalculateGradient(lo, hi, full, x0)
    assert np.array_equal(g, g_ref)
    c.stop_optimization_ = True
    with pytest.raises(ForcedStop):
        c.EvaluateObjective(xf, x0)
    assert c.stop_optimization_ is False        # the flag is consumed (core_opt.cpp:52)
    assert c.EvaluateObjective(xf, x0) == f


def test_unsorted_chain_and_out_of_range_samples(oracle_port):
    """Control-point times are parameters: when a candidate moves a key past its neighbour the reference's
    first-match-in-list-order segment search decides (curve.cpp:100-128)."""
    scene = load_scene("stick-figure")
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1)
    c = Core(scene_path("stick-figure"))
    x0 = c.GetDescriptor()
    x = x0.copy()
    x[0] = 11.5        # first key of the first keyed channel jumps over the second one (t = 10)
    x[5 * 7] = 2.0     # another channel: a key moved before the first one
    lo, hi = c.window
    f = c.CalculateFullObjective(lo, hi, x, x0)
    fo, _ = o.full_objective(x, x0)
    assert abs(f - fo) <= TOL * abs(fo)
    # a window far outside every key: all curves clamp to their end values -> zero torque, zero cost
    f_far = c.CalculateFullObjective(100.0, 120.0, x0, x0)
    o.update_original(x0)
    fo_far, _ = o.full_objective(x0, x0, window=(100.0, 120.0))
    assert abs(f_far - fo_far) <= 1e-12 and abs(f_far) < 1e-20


def test_small_windows_and_boundary_stencils(oracle_port):
    """N_s = 3, 4, 5, 6 exercise every branch of the one-sided / 3-point / 5-point stencils (core_rbdl.cpp:28-95)."""
    scene = load_scene("fox-tail")
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1)
    c = Core(scene_path("fox-tail"))
    x0 = c.GetDescriptor()
    for hi in (10.45, 10.65, 10.85, 11.05, 35.0):
        f = c.CalculateFullObjective(10.0, hi, x0, x0)
        o.update_original(x0)   # the reference's OriginalMotionCache is keyed on x_original only: reset it per window
        fo, _ = o.full_objective(x0, x0, window=(10.0, hi))
        assert abs(f - fo) <= TOL * abs(fo) + 1e-15, (hi, f, fo)
    with pytest.raises(EngineError):
        c.CalculateFullObjective(10.0, 10.3, x0, x0)      # 2 samples: the reference reads out of bounds there


def test_config4_batch_properties(oracle_port):
    """BASELINE config 4 shape (stick-figure x 240 frames, D = 780, N_s = 1216) at a reduced candidate count
    for the oracle comparison plus size-independent properties at the full 4096."""
    scene = synthetic.stick_240(load_scene("stick-figure"))
    c = Core(scene)
    x0 = c.GetDescriptor()
    lo, hi = c.window
    assert c.D == 780
    B = 4096
    X = synthetic.perturbed_candidates(x0, B)
    f, terms = c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)
    assert np.isfinite(f).all() and (f > 0).all()
    assert np.array_equal(f, terms[:, 0] + terms[:, 1] + terms[:, 2])
    # permutation invariance / idempotence: rows are independent and deterministic
    perm = np.random.default_rng(0).permutation(B)
    f2 = c.CalculateFullObjectiveBatch(lo, hi, X[perm], x0)
    assert np.array_equal(f2, f[perm])
    # D_param is an exact quadratic form of x - x0 once weights are set; D_traj >= 0
    c.UpdateGlobalWeight(12.5)
    f3, t3 = c.CalculateFullObjectiveBatch(lo, hi, X[:64], x0, want_terms=True)
    assert np.allclose(t3[:, 1], 12.5 * np.sum((X[:64] - x0) ** 2, axis=1), rtol=1e-13)
    assert (t3[:, 2] >= 0).all() and np.allclose(t3[:, 0], terms[:64, 0], rtol=1e-15, atol=0)
    f_same, t_same = c.CalculateFullObjective(lo, hi, x0, x0, want_terms=True)
    assert t_same[1] == 0.0 and abs(t_same[2]) <= 1e-20
    # oracle on a few rows
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1); o.set_weight_trajectory(12.5)
    for b in (0, 17, 63):
        fo, _ = o.full_objective(X[b], x0)
        assert abs(f3[b] - fo) <= TOL * abs(fo)
    # sharded gradient columns == monolithic gradient (the multi-GPU decomposition on one device)
    eng = c.engine()
    idx = c.free_indices()
    f0, g, fpm = eng.eval_grad(X[1], idx, 1e-4)
    P = 2 * idx.size + 1
    cols = np.zeros(P)
    d_out = eng.device_alloc(8 * P)
    for r in range(3):
        b, n = shard_range(P, r, 3)
        eng.eval_grad_columns(X[1], idx, 1e-4, b, n, d_out, sync=True)
        cols[b:b + n] = eng.d2h(d_out, (n,))
    eng.device_free(d_out)
    f0s, gs = grad_from_columns(cols, 1e-4)
    assert f0s == f0 and np.array_equal(gs, g)


def test_config5_tree_matches_oracle(oracle_port):
    """BASELINE config 5 skeleton (64-joint binary tree, dense keys) on a short window."""
    scene = synthetic.tree_scene(n_joints=64, n_frames=60)
    c = Core(scene)
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1)
    x0 = c.GetDescriptor()
    X = synthetic.perturbed_candidates(x0, 3)
    lo, hi = c.window
    c.UpdateGlobalWeight(1.0); o.set_weight_trajectory(1.0)
    f = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
    for b in range(3):
        fo, _ = o.full_objective(X[b], x0)
        assert abs(f[b] - fo) <= TOL * abs(fo), (f[b], fo)


def test_counters_report_executed_curve_and_ik_work():
    c = Core(scene_path("conductor"))
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    eng.set_counting(True)
    eng.eval(x0)
    k = eng.counters()
    eng.set_counting(False)
    assert k["ik_solves"] == 220                        # one solve per time sample (N_s = 220)
    assert 12 * 220 <= k["ik_iters"] <= 20 * 220
    assert 0 < k["bezier_evals"] <= 220 * 4 and k["newton_iters"] >= k["bezier_evals"]
    assert k["launches"] > 0


def test_hand_scheduled_division_is_ieee_exact():
    """div_rn_fast (the curve solver's branch-free f / f') == the compiler's IEEE division bit for bit wherever its own
    validity test passes; operands of the size Newton produces must all pass it."""
    c = Core(scene_path("stick-figure"))
    eng = c.engine()
    rng = np.random.default_rng(7)
    n = 1 << 20
    # Newton-like operands, a wide log-uniform sweep, and edge cases (zeros, denormals, infinities, NaN, huge ratios)
    num = np.concatenate([rng.normal(size=n) * 10.0 ** rng.uniform(-12, 2, n), rng.normal(size=n) * 10.0 ** rng.uniform(-300, 300, n),
                          np.array([0.0, -0.0, 1.0, 5e-324, 1e-310, np.inf, -np.inf, np.nan, 1e308, 1e-308, 3.0, 1.0])])
    den = np.concatenate([rng.normal(size=n) * 10.0 ** rng.uniform(-3, 3, n), rng.normal(size=n) * 10.0 ** rng.uniform(-300, 300, n),
                          np.array([1.0, 1.0, 0.0, 1.0, 3.0, 1.0, np.inf, 1.0, 1e-308, 1e308, 5e-324, 3.0])])
    qf, qr, ok = eng.selftest_division(num, den)
    sel = ok != 0
    assert np.array_equal(qf[sel].view(np.uint64), qr[sel].view(np.uint64))
    assert sel[:n].all()                       # the Newton-sized operands never need the full-range routine
    with np.errstate(all="ignore"):
        host = num / den
    fin = np.isfinite(host)
    assert np.array_equal(qr[fin].view(np.uint64), host[fin].view(np.uint64))   # and the device reference is IEEE


def _forced(monkeypatch, mode, **env):
    """A fresh Core whose engine is forced onto (OMO_JIT=1) or off (OMO_JIT=0) the scene-specialised kernel."""
    monkeypatch.setenv("OMO_JIT", mode)
    for k, v in env.items():
        monkeypatch.setenv(k, v)


@pytest.mark.parametrize("form", ["unroll", "loop"])
def test_specialised_kernel_paths_match_table_driven_kernel(monkeypatch, oracle_port, form):
    """The scene-specialised kernel (staged segment windows, structural zeros, compact stack; both code forms) against the
    table-driven kernel and the oracle on the cases that leave its fast path: a candidate with an unsorted chain, a
    sample window that starts before the first key and ends after the last, keys denser than the staged window
    (OMO_SEGW segments per 128 samples), a single-key channel, trajectory weight (FK instance) and an IK scene."""
    cases = []
    # (a) stick-figure: unsorted chain + out-of-range samples + FK instance
    scene = load_scene("stick-figure")
    x0 = Core(scene).GetDescriptor()
    xa = x0.copy(); xa[0] = 11.5; xa[5 * 7] = 2.0
    rng = np.random.default_rng(3)
    xb = x0.copy(); xb[2::5] += 0.2 * (rng.random(xb.size // 5) - 0.5)
    cases.append((scene, np.stack([xa, xb, x0]), 12.5, (-20.0, 90.0)))
    # (b) dense keys: a key every frame = 5 samples per segment, 25 segments per block -> every thread beyond the window
    dense = synthetic.tree_scene(n_joints=6, n_frames=40, key_every=1)
    xd = Core(dense).GetDescriptor()
    cases.append((dense, synthetic.perturbed_candidates(xd, 3), 0.0, None))
    # (c) a single-key channel next to keyed ones
    single = load_scene("fox-tail")
    def find(node, name):
        if node["name"] == name:
            return node
        for ch in node.get("children", []):
            hit = find(ch, name)
            if hit:
                return hit
        return None
    find(single["joints"], "tail_root")["rot_x"] = {"var": 0.0, "curve": [[12.0, 0.3, 0.1, -1.0, 1.0]]}
    xs = Core(single).GetDescriptor()
    cases.append((single, np.stack([xs, xs * 1.0]), 1.0, None))
    # (d) IK scene
    cond = load_scene("conductor")
    xc = Core(cond).GetDescriptor()
    xc2 = xc.copy(); xc2[2::5] += 0.1 * (rng.random(xc2.size // 5) - 0.5)
    cases.append((cond, np.stack([xc, xc2]), 12.5, None))

    for scene, X, w_traj, window in cases:
        res = {}
        for mode in ("0", "1"):
            _forced(monkeypatch, mode, OMO_JIT_FORM=form)
            c = Core(scene)
            c.UpdateGlobalWeight(w_traj)
            lo, hi = window if window else c.window
            x0 = c.GetDescriptor()
            res[mode] = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
            if mode == "1":
                assert c.engine().jit_info()["launches"] > 0, c.engine().jit_info()["log"]
        assert np.allclose(res["1"], res["0"], rtol=1e-13, atol=0), (res["1"], res["0"])
        o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1); o.set_weight_trajectory(w_traj)
        x0 = Core(scene).GetDescriptor()
        o.update_original(x0)
        fo, _ = o.full_objective(X[0], x0, window=window) if window else o.full_objective(X[0], x0)
        assert abs(res["1"][0] - fo) <= TOL * abs(fo) + 1e-15, (res["1"][0], fo)


@pytest.mark.parametrize("name", ["stick-figure", "conductor", "stick240"])
def test_sparse_gradient_columns_equal_dense_evaluation(monkeypatch, name):
    """Finite-difference columns evaluated only on the tiles their control point can reach (include/omo.h:
    omo_grad_sparsity) give bit for bit the f(x +- h e_k) of evaluating every column on every sample."""
    scene = synthetic.stick_240(load_scene("stick-figure")) if name == "stick240" else load_scene(name)
    res = {}
    monkeypatch.setenv("OMO_JIT", "1")     # both ways through the scene-specialised kernel (sparse launches always use it)
    for dense in ("1", None):
        if dense:
            monkeypatch.setenv("OMO_GRAD_DENSE", dense)
        else:
            monkeypatch.delenv("OMO_GRAD_DENSE", raising=False)
        c = Core(scene)
        c.UpdateGlobalWeight(12.5)
        x0 = c.GetDescriptor()
        rng = np.random.default_rng(11)
        x = x0.copy(); x[2::5] += 0.2 * (rng.random(x.size // 5) - 0.5)
        lo, hi = c.window
        f0, g, fpm = c.CalculateGradient(lo, hi, x, x0)
        res[dense] = (f0, g, fpm, c.engine().grad_sparsity())
    assert res[None][0] == res["1"][0]
    assert np.array_equal(res[None][2], res["1"][2]) and np.array_equal(res[None][1], res["1"][1])
    executed, dense_items = res[None][3]
    assert 0 < executed
    if name == "stick240":
        assert executed < dense_items / 3      # 10 tiles per column, one or two of them within reach of a control point


@pytest.mark.parametrize("name", SCENES)
def test_single_candidate_graph_replay_equals_plain_launches(monkeypatch, name):
    """omo_eval with B = 1 (the NLopt callback's call, core_opt.cpp:42-57) replays a captured CUDA graph from its third
    call on. Same kernels, same order: bit-identical to the plain launch sequence (OMO_GRAPH=0), across weight, kernel,
    x_original and window changes made between calls (what the dialogs do mid-solve, sliderdialog.cpp:71-76)."""
    monkeypatch.setenv("OMO_SELF", "0")      # the one-launch path (tested below) would pre-empt the graph
    def run(graph):
        monkeypatch.setenv("OMO_GRAPH", "1" if graph else "0")
        c = Core(scene_path(name))
        x0 = c.GetDescriptor()
        idx = c.free_indices()
        lo, hi = c.window
        out = []
        r = np.random.default_rng(11)
        for step in range(12):
            x = x0.copy()
            x[idx] += 0.01 * r.standard_normal(idx.size)
            if step == 4:
                c.UpdateGlobalWeight(12.5)                 # w_traj: 0 -> 12.5 switches the FK instance of the kernel
            if step == 7:
                c.set_kernels([(1.0, 0.5 * (lo + hi), 2.0)])
            if step == 9:
                hi = hi - 3.0                              # another window: other grid, other buffers
            f, terms = c.CalculateFullObjective(lo, hi, x, x0 if step < 10 else x, want_terms=True)
            out.append((f,) + tuple(terms))
        return np.array(out), c.engine().graph_replays(), c.engine().counters()["launches"]

    plain, n_plain, l_plain = run(False)
    graph, n_graph, l_graph = run(True)
    assert n_plain == 0
    assert n_graph >= 4, n_graph
    assert l_graph == l_plain           # replays are counted as the kernels they launch
    assert np.array_equal(plain, graph)


@pytest.mark.parametrize("root,nb", [("LeftForeArm", 1), ("LeftArm", 2), ("LeftShoulder", 3), ("Spine", 4)])
def test_ik_chain_lengths_match_oracle(monkeypatch, oracle_port, root, nb):
    """IkHandler chains of 1-4 bodies (ikhandler.cpp:166-188: one EulerZYX body per joint between root and end effector):
    the IK solve is specialised on the chain length for 1, 2 and 3 bodies and generic beyond (ik_solve_n, omo_kernels.cuh).
    The conductor scene with its handle's root moved up and down the left arm, table-driven and scene-specialised kernel."""
    import copy
    scene = copy.deepcopy(load_scene("conductor"))
    scene["ik_handlers"][0]["root_name"] = root
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1); o.set_weight_trajectory(12.5)
    rng = np.random.default_rng(nb)
    res = {}
    for mode in ("0", "1"):
        _forced(monkeypatch, mode)
        c = Core(scene)
        c.UpdateGlobalWeight(12.5)
        x0 = c.GetDescriptor()
        X = np.stack([x0, x0])
        X[1, 2::5] += 0.1 * (np.random.default_rng(nb).random(x0.size // 5) - 0.5)
        lo, hi = c.window
        c.engine(synced_window=(lo, hi)).set_counting(True)
        res[mode] = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
        k = c.engine().counters()
        assert k["ik_solves"] == 2 * c.engine().num_samples, k
        assert k["ik_iters"] >= k["ik_solves"]
        if mode == "1":
            assert c.engine().jit_info()["launches"] > 0, c.engine().jit_info()["log"]
    assert np.allclose(res["1"], res["0"], rtol=1e-13, atol=0), (res["1"], res["0"])
    for b in range(2):
        fo, _ = o.full_objective(X[b], x0)
        assert abs(res["0"][b] - fo) <= TOL * abs(fo), (root, b, res["0"][b], fo)


def _rechain(scene):
    """The joints of a tree_scene re-linked as ONE chain (the deepest tree a joint count allows)."""
    nodes = list(synthetic.iter_joints(scene["joints"]))
    for n in nodes:
        n["children"] = []
    for a, b in zip(nodes, nodes[1:]):
        a["children"] = [b]
    return scene


@pytest.mark.parametrize("shape", ["binary", "chain"])
def test_maximum_joint_count_and_empty_inputs(oracle_port, shape):
    """Edge sizes of the C ABI (include/omo.h): OMO_MAX_JOINTS = 128 joints as a binary tree (widest sibling stack) and as a
    single chain (deepest recursion, 127 nested bodies), against the oracle; one joint more is refused, not truncated;
    an empty batch is a no-op; a scene without any keyed curve (D = 0) evaluates to its constant-motion cost."""
    scene = synthetic.tree_scene(n_joints=128, n_frames=12, key_every=4)
    if shape == "chain":
        scene = _rechain(scene)
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1); o.set_weight_trajectory(3.0)
    c = Core(scene)
    c.UpdateGlobalWeight(3.0)
    x0 = c.GetDescriptor()
    X = synthetic.perturbed_candidates(x0, 3)
    lo, hi = c.window
    f = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
    for b in range(3):
        fo, _ = o.full_objective(X[b], x0)
        assert abs(f[b] - fo) <= TOL * abs(fo), (shape, b, f[b], fo)
    assert c.CalculateFullObjectiveBatch(lo, hi, X[:0], x0).size == 0

    too_many = synthetic.tree_scene(n_joints=129, n_frames=12, key_every=4)
    if shape == "chain":
        too_many = _rechain(too_many)
    with pytest.raises(EngineError, match="OMO_MAX_JOINTS"):
        big = Core(too_many)
        big.CalculateFullObjective(*big.window, big.GetDescriptor(), big.GetDescriptor())

    if shape == "binary":
        still = synthetic.tree_scene(n_joints=6, n_frames=12, key_every=4)
        for n in synthetic.iter_joints(still["joints"]):
            for k, v in (("rot_x", 0.3), ("rot_y", -0.2), ("rot_z", 0.1)):
                n[k] = {"var": v, "curve": []}
        cs = Core(still)
        xs = cs.GetDescriptor()
        assert xs.size == 0
        os_ = oracle_port.Oracle(still, "port")
        fs = cs.CalculateFullObjective(*cs.window, xs, xs)
        fo, _ = os_.full_objective(xs, xs)
        assert abs(fs - fo) <= 1e-9 * abs(fo) + 1e-18, (fs, fo)


def test_free_parameter_batch_equals_expanded_batch():
    """omo_eval_free (candidates as their F free entries, expanded on the device like EvaluateObjective's revert map,
    core_opt.cpp:99-121) against omo_eval on the host-expanded rows: bit for bit, small and chunked batches."""
    scene = synthetic.stick_240(load_scene("stick-figure"))
    c = Core(scene)
    c.UpdateGlobalWeight(2.0)
    x0 = c.GetDescriptor()
    idx = c.free_indices()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    for B in (1, 7, 1500):                       # 1500 x 468 doubles = 5.6 MB -> the pipelined, chunked path
        X = synthetic.perturbed_candidates(x0, B)
        f_full, t_full = eng.eval(X, want_terms=True)
        f_free, t_free = eng.eval_free(x0, idx, X[:, idx], want_terms=True)
        assert np.array_equal(f_full, f_free) and np.array_equal(t_full, t_free), B
    with pytest.raises(EngineError):
        eng.eval_free(x0, np.array([x0.size], dtype=np.int32), np.zeros((1, 1)))


# ---- full-size configurations against the oracle ("close the untested configs") -------------------------------

def test_config5_full_size_tree_matches_oracle(oracle_port):
    """BASELINE config 5 at its FULL size -- 64-joint binary tree x 2000 frames, 10015 samples, D = 189 945 -- through the
    scene-specialised kernel, against the oracle on 8 of the candidates (one objective of this size takes the CPU ~1 s)."""
    scene = synthetic.tree_scene(64, 2000)
    c = Core(scene)
    x0 = c.GetDescriptor()
    assert c.D == 189945
    lo, hi = c.window
    X = synthetic.perturbed_candidates(x0, 64)
    f = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
    assert c.engine().num_samples == 10015
    assert c.engine().jit_info()["launches"] > 0, c.engine().jit_info()["log"]
    o = oracle_port.Oracle(scene, "port")
    sec, fo = o.bench_batch(X[:8], 1, 0)
    assert np.max(np.abs(f[:8] - fo) / np.abs(fo)) <= TOL, np.max(np.abs(f[:8] - fo) / np.abs(fo))
    # the other rows: finite, positive, and independent of their position in the batch
    assert np.isfinite(f).all() and (f > 0).all()
    f2 = c.CalculateFullObjectiveBatch(lo, hi, X[::-1].copy(), x0)
    assert np.array_equal(f2[::-1], f)


@pytest.mark.parametrize("name", ["stick240", "conductor", "stick-figure", "fox-tail"])
def test_gradient_matches_oracle_gradient(oracle_port, name):
    """The finite-difference gradient (sparse columns, scene-specialised kernel) against Oracle.gradient -- the same quotient
    formed from the CPU objective -- at the sizes of BASELINE configs 1, 3 and 4: 937 columns for stick240, 109 with the IK
    handle, two Gaussian kernels and w_traj = 12.5 for the conductor."""
    scene = synthetic.stick_240(load_scene("stick-figure")) if name == "stick240" else load_scene(name)
    c = Core(scene)
    o = oracle_port.Oracle(scene, "port")
    if name == "conductor":
        ks = [(1.0, 20.0, 2.6), (1.0, 10.0, 1.0)]
        c.set_kernels(ks); o.set_kernels(ks)
    if name != "fox-tail":
        c.UpdateGlobalWeight(12.5); o.set_weight_trajectory(12.5)
    x0 = c.GetDescriptor()
    x = synthetic.perturbed_candidates(x0, 1, first=3)[0]
    lo, hi = c.window
    assert np.array_equal(c.free_indices(), o.free_indices())
    f0, g, fpm = c.CalculateGradient(lo, hi, x, x0, h=1e-4)
    o.update_original(x0)
    fo, go, fpmo = o.gradient(x, h=1e-4)
    assert g.size == {"stick240": 468, "conductor": 54, "stick-figure": 78, "fox-tail": 27}[name]
    assert abs(f0 - fo) <= TOL * abs(fo)
    assert np.max(np.abs(fpm - fpmo) / np.abs(fpmo)) <= TOL, np.max(np.abs(fpm - fpmo) / np.abs(fpmo))
    assert np.max(np.abs(g - go)) <= gradient_bound(go, fo, 1e-4), (np.max(np.abs(g - go)), gradient_bound(go, fo, 1e-4))


@pytest.mark.parametrize("root,nb", [("LeftForeArm", 1), ("LeftArm", 2), ("LeftShoulder", 3), ("Spine", 4)])
def test_randomised_ik_sweep_values_and_iteration_counts(oracle_port, root, nb):
    """64 randomly perturbed conductor candidates per chain length x 220 samples: f within 1e-9 AND the executed work -- damped
    least-squares iterations (both exits of RBDL's loop, Kinematics.cc:660,685, compare against 1e-6) -- equal to the oracle's
    counters candidate by candidate. A sample whose |e| or |dq| lands within
    rounding of the threshold would leave the loop one iteration apart from the CPU and show up here."""
    import copy
    scene = copy.deepcopy(load_scene("conductor"))
    scene["ik_handlers"][0]["root_name"] = root
    o = oracle_port.Oracle(scene, "port"); o.set_sampling_concurrency(1); o.set_weight_trajectory(12.5)
    c = Core(scene)
    c.UpdateGlobalWeight(12.5)
    x0 = c.GetDescriptor()
    B = 64
    X = synthetic.perturbed_candidates(x0, B, seed_base=0x1C00 + nb)
    # move the handle's target values too (theta / t_b / t_f only is what perturbed_candidates does): +-3 cm on every key value
    rng = np.random.default_rng(100 + nb)
    ik_first = 5 * (x0.size // 5 - 5)
    X[:, ik_first + 1::5] += 0.03 * (rng.random((B, 5)) - 0.5)
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    f = c.CalculateFullObjectiveBatch(lo, hi, X, x0)
    o.update_original(x0)
    o.full_objective(x0, x0)                    # fills the oracle's OriginalMotionCache so that the counts below are one SetStates each
    bad = []
    for b in range(B):
        oracle_port.counters("port", reset=True)
        fo, _ = o.full_objective(X[b], x0)
        ko = oracle_port.counters("port")
        assert abs(f[b] - fo) <= TOL * abs(fo), (root, b, f[b], fo)
        eng.set_counting(True)
        fb = eng.eval(X[b])
        kg = eng.counters()
        eng.set_counting(False)
        assert fb[0] == f[b]
        # (Bezier / Newton counts are not comparable: the reference re-evaluates every ancestor's curves for each joint's world
        # affine, joint.cpp:76-83, the kernel samples each curve once per time sample)
        if (kg["ik_solves"], kg["ik_iters"]) != (ko["ik_solves"], ko["ik_iters"]):
            bad.append((b, kg["ik_solves"], kg["ik_iters"], ko["ik_solves"], ko["ik_iters"]))
    assert not bad, bad[:6]


@pytest.mark.parametrize("name", ["stick-figure", "fox-tail", "conductor", "stick240"])
def test_one_launch_path_is_bit_identical_to_the_general_path(monkeypatch, name):
    """Single candidates (every BOBYQA callback, core_opt.cpp:145) and small finite-difference gradients run as ONE launch of the
    self-preparing instance of the scene-specialised kernel: x as a kernel parameter, control points prepared per block in shared
    memory, the last block reduces, f written to mapped host memory. Same arithmetic, same order as perturb -> prepare ->
    objective -> finish: bit-identical f, terms, f(x +- h e_k) and gradient -- with weights, kernels, a cyclic curve, an unsorted
    chain and a window that some control points cannot reach."""
    scene = synthetic.stick_240(load_scene("stick-figure")) if name == "stick240" else load_scene(name)
    monkeypatch.setenv("OMO_JIT", "1")        # the general path through the scene-specialised kernel too
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("OMO_SELF", mode)
        c = Core(scene)
        x0 = c.GetDescriptor()
        idx = c.free_indices()
        lo, hi = c.window
        rng = np.random.default_rng(21)
        out = []
        for step in range(6):
            x = x0.copy()
            x[idx] += 0.02 * rng.standard_normal(idx.size)
            if step == 1:
                c.UpdateGlobalWeight(12.5)
            if step == 2:
                c.set_kernels([(1.0, 0.5 * (lo + hi), 2.0)])
            if step == 3:
                c.set_cyclic(True)
            if step == 4:
                c.set_cyclic(False)
                x[0] = x[5] + 0.25                  # first key of the first keyed channel jumps over the second: unsorted chain
            if step == 5:
                hi = lo + 0.45 * (hi - lo)          # a shorter window: the later control points are out of its reach
            f, terms = c.CalculateFullObjective(lo, hi, x, x0, want_terms=True)
            f0, g, fpm = c.CalculateGradient(lo, hi, x, x0)
            out.append((f, tuple(terms), f0, g.copy(), fpm.copy()))
        res[mode] = (out, c.engine().jit_info()["launches"])
    for a, b in zip(res["0"][0], res["1"][0]):
        assert a[0] == b[0] and a[1] == b[1] and a[2] == b[2]
        assert np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])
        assert a[0] == a[2]                          # f of the single call == f0 of the gradient call
    assert res["1"][1] > 0


@pytest.mark.parametrize("name", ["stick240", "tree24"])
def test_sorted_curve_work_lists_change_no_bit(monkeypatch, name):
    """Large launches of the scene-specialised kernel solve each tile's (channel, sample) curve items in the order of the Newton
    iterations x_original's motion needs for them (OmoEvalParams::perm, omo.cu: build_perm), so that a warp's lockstep triple holds
    curves of similar cost. Which lane solves which curve must not change a bit of any value -- against the per-sample order
    (OMO_SORTED_CURVES=0), with an x_original that is then replaced, a candidate with an unsorted chain in the batch, and w_traj."""
    scene = synthetic.stick_240(load_scene("stick-figure")) if name == "stick240" else synthetic.tree_scene(n_joints=24, n_frames=400)
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("OMO_SORTED_CURVES", mode)
        c = Core(scene)
        x0 = c.GetDescriptor()
        lo, hi = c.window
        B = 1200 if name == "stick240" else 400
        X = synthetic.perturbed_candidates(x0, B)
        X[7, 0] = X[7, 5] + 0.25                       # one candidate's first chain becomes unsorted
        out = [c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)]
        c.UpdateGlobalWeight(2.5)
        out.append(c.CalculateFullObjectiveBatch(lo, hi, X, X[3], want_terms=True))     # another x_original: the lists are rebuilt
        res[mode] = out
    for a, b in zip(res["0"], res["1"]):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("name", ["tree24", "stick240"])
def test_threshold_segment_search_changes_no_bit(monkeypatch, name):
    """Scenes without sorted work lists find a curve's segment through per-channel thresholds in sample-index space (plan_build /
    eval_channels_planned: "key time <= t" as "slot >= threshold", a byte-wise SIMD compare); OMO_PLAN=0 keeps the five FP64 compares per
    curve and sample (eval_channels_staged). Same segment, same record, same value -- with an unsorted chain in the batch, key times
    outside the window's staged records and w_traj. (stick240 runs its per-sample path here: sorted lists off.)"""
    scene = synthetic.stick_240(load_scene("stick-figure")) if name == "stick240" else synthetic.tree_scene(n_joints=24, n_frames=400)
    monkeypatch.setenv("OMO_SORTED_CURVES", "0")
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("OMO_PLAN", mode)
        c = Core(scene)
        x0 = c.GetDescriptor()
        lo, hi = c.window
        B = 1200 if name == "stick240" else 400
        X = synthetic.perturbed_candidates(x0, B)
        X[7, 0] = X[7, 5] + 0.25                       # one candidate's first chain becomes unsorted
        X[9, 5::5] = X[9, 5::5] + 1e-3                 # another one's key times moved off the sample grid
        out = [c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)]
        c.UpdateGlobalWeight(2.5)
        out.append(c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True))
        assert c.engine().jit_info()["launches"] > 0
        res[mode] = out
    for a, b in zip(res["0"], res["1"]):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_curve_group_forms_change_no_bit(monkeypatch):
    """A scene that takes no sorted work lists gets its curve groups as curve_group_n<K, REM> (omo_jit.hpp curve_k): the per-sample
    triples without the sorted form compiled beside them. OMO_CURVE_K=0 restores curve_group<REM>, OMO_CURVE_K=4 solves four curves per
    lockstep round (remainder groups included): which curves share a round changes nothing in a curve's own arithmetic. With an
    unsorted chain, key times off the sample grid and w_traj, as above."""
    scene = synthetic.tree_scene(n_joints=24, n_frames=400)
    res = {}
    for mode in ("0", None, "4"):
        if mode is None:
            monkeypatch.delenv("OMO_CURVE_K", raising=False)
        else:
            monkeypatch.setenv("OMO_CURVE_K", mode)
        c = Core(scene)
        x0 = c.GetDescriptor()
        lo, hi = c.window
        X = synthetic.perturbed_candidates(x0, 400)
        X[7, 0] = X[7, 5] + 0.25
        X[9, 5::5] = X[9, 5::5] + 1e-3
        out = [c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)]
        c.UpdateGlobalWeight(2.5)
        out.append(c.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True))
        assert c.engine().jit_info()["launches"] > 0
        res[mode] = out
    for mode in (None, "4"):
        for a, b in zip(res["0"], res[mode]):
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])

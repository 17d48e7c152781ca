"""The multi-device engine (include/omo.h: omo_create_multi ...) and the thread safety of the headless Core.

One host thread, several GPUs: batches sharded over candidates, gradients over their 2F+1 finite-difference columns
(SURVEY.md section 8(e)); results must be bit-identical to the single-device entry points. Runs on however many GPUs the box
has (1 on the default test box: the worker / shard / gather plumbing is the same, the slices just land on one device)."""
import ctypes
import os
import threading

import numpy as np
import pytest

from conftest import load_scene, scene_path

from optimo_b200 import synthetic
from optimo_b200.core import Core, EngineError, MultiEngine

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        cudart = ctypes.CDLL("libcudart.so")
    except OSError:
        cudart = ctypes.CDLL("/usr/local/cuda/lib64/libcudart.so")
    n = ctypes.c_int(0)
    return n.value if cudart.cudaGetDeviceCount(ctypes.byref(n)) == 0 else 0


@pytest.mark.parametrize("gather", ["host", "nccl"])
def test_multi_engine_is_bit_identical_to_one_device(gather):
    """Core.SetDevices([...]) routes CalculateFullObjectiveBatch / CalculateGradient through omo_multi_eval / omo_multi_eval_grad
    (candidates and columns split by omo_shard_range, gathered by per-device D2H or by ncclAllGather): same bits as one device."""
    n = _n_gpus()
    devices = list(range(max(n, 1)))
    scene = synthetic.stick_240(load_scene("stick-figure"))
    one = Core(scene, device=0)
    many = Core(scene, devices=devices if len(devices) > 1 else [0])
    if len(devices) < 2:
        # a one-entry list means "one device": the multi engine must not exist
        with pytest.raises(EngineError):
            many.multi_engine()
        many = None
    x0 = one.GetDescriptor()
    lo, hi = one.window
    X = synthetic.perturbed_candidates(x0, 301)          # not a multiple of any device count
    one.UpdateGlobalWeight(12.5)
    f1, t1 = one.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)
    g1 = one.CalculateGradient(lo, hi, X[2], x0)
    if many is not None:
        many.UpdateGlobalWeight(12.5)
        m = many.multi_engine()
        assert m.num_devices == len(devices) and m.devices() == devices
        m.set_gather(MultiEngine.GATHER_NCCL if gather == "nccl" else MultiEngine.GATHER_HOST)
        fm, tm = many.CalculateFullObjectiveBatch(lo, hi, X, x0, want_terms=True)
        gm = many.CalculateGradient(lo, hi, X[2], x0)
        assert np.array_equal(f1, fm) and np.array_equal(t1, tm)
        assert g1[0] == gm[0] and np.array_equal(g1[1], gm[1]) and np.array_equal(g1[2], gm[2])
        # single evaluations run on the first device of the list, same bits again
        assert many.CalculateFullObjective(lo, hi, X[7], x0) == f1[7]


def test_multi_engine_raw_abi_on_available_devices():
    """omo_create_multi called directly: every visible device (n_dev <= 0), slices of a batch smaller than the device count,
    an empty batch, and the error paths (bad device list -> OMO_BAD_ARG with a message)."""
    from optimo_b200.core import _omo, _dp, _ip
    n = max(_n_gpus(), 1)
    scene = load_scene("fox-tail")
    c = Core(scene, device=0)
    x0 = c.GetDescriptor()
    lo, hi = c.window
    f_ref = c.CalculateFullObjectiveBatch(lo, hi, np.stack([x0, x0 + 0.01]), x0)
    idx = c.free_indices() if c.free_indices().size else np.arange(2, x0.size, 5, dtype=np.int32)
    g_ref = c.CalculateGradient(lo, hi, x0, x0, free_idx=idx)

    # the same scene through a Core that owns a multi engine over all devices (two slots on a 1-GPU box are not allowed:
    # device ids must be distinct, so there it is the error path that is exercised)
    if n >= 2:
        cm = Core(scene, devices=list(range(n)))
        fm = cm.CalculateFullObjectiveBatch(lo, hi, np.stack([x0, x0 + 0.01]), x0)
        assert np.array_equal(fm, f_ref)
        gm = cm.CalculateGradient(lo, hi, x0, x0, free_idx=idx)
        assert gm[0] == g_ref[0] and np.array_equal(gm[1], g_ref[1])
        assert cm.CalculateFullObjectiveBatch(lo, hi, np.zeros((0, x0.size)), x0).size == 0
        # one candidate over n devices: n - 1 empty slices; same bits as the one-candidate call on one device
        assert cm.CalculateFullObjectiveBatch(lo, hi, x0[None, :], x0)[0] == c.CalculateFullObjectiveBatch(lo, hi, x0[None, :], x0)[0]
    with pytest.raises(EngineError, match="distinct"):
        bad = Core(scene, devices=[0, 0])
        bad.CalculateFullObjectiveBatch(lo, hi, x0[None, :], x0)
    with pytest.raises(EngineError):
        bad = Core(scene, devices=[0, 4096])
        bad.CalculateFullObjectiveBatch(lo, hi, x0[None, :], x0)


def test_ui_thread_queries_during_a_solve_are_serialised():
    """the reference's UI thread calls GetOptimizationCache at 30 Hz and CalculateTorques per redraw while the
    solver thread evaluates (core.h:132-143, mainwindow.cpp:228-235). Every engine sequence of the headless Core holds its engine
    mutex, so the queries interleave with a running PerformOptimization without corrupting it: same iterates and same result as
    an undisturbed solve, and every cache the UI thread saw is a consistent evaluation."""
    def solve(disturb):
        c = Core(scene_path("fox-tail"))
        c.select_items(["tail_root", "tail_center"])
        c.record_solve_trace(True)
        seen, errors, stop = [], [], threading.Event()

        def ui():
            try:
                while not stop.is_set():
                    cache = c.GetOptimizationCache()
                    tau = c.CalculateTorques([3.0, 7.5, 20.0])
                    seen.append((cache["time"].size, float(np.sum(cache["local_cost"])), float(np.abs(tau).max())))
            except Exception as e:   # noqa: BLE001 -- reported below
                errors.append(e)

        th = threading.Thread(target=ui)
        if disturb:
            th.start()
        f, evals = c.PerformOptimization(False)
        stop.set()
        if disturb:
            th.join()
        assert not errors, errors
        X, F = c.solve_trace()
        return f, evals, X, F, seen

    f0, e0, X0, F0, _ = solve(False)
    f1, e1, X1, F1, seen = solve(True)
    assert (f0, e0) == (f1, e1)
    assert np.array_equal(X0, X1) and np.array_equal(F0, F1)
    assert len(seen) > 3
    assert all(n in (235, 245) and np.isfinite(s) and np.isfinite(t) for n, s, t in seen)   # 235: before the first evaluation (core.h:136-139)


def test_peer_gather_between_two_processes_equals_the_whole_batch(tmp_path):
    """omo_peer_gather_* (include/omo.h): two PROCESSES evaluate the halves of one batch; the kernel that forms f stores it into both
    ranks' gathered arrays through cudaIpc-mapped peer memory and waits for the other rank's flag. Every rank must end up with the
    bits a single device computes for the whole batch -- on both halves of the double-buffered array, and with a rank that has no
    candidates in a call. (One GPU: both ranks share device 0 and time-slice; two or more: one device each.)"""
    import subprocess
    import sys
    n = _n_gpus()
    B, world = 257, 2
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "peer_gather_worker.py")
    procs = [subprocess.Popen([sys.executable, worker, str(tmp_path), str(r), str(world), str(r if n >= world else 0), str(B)],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(world)]
    logs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        logs.append(out.decode(errors="replace")[-2000:])
    assert all(p.returncode == 0 for p in procs), logs
    c = Core(synthetic.stick_240(load_scene("stick-figure")), device=0)
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    got = [np.load(os.path.join(str(tmp_path), f"out{r}.npy")) for r in range(world)]
    for call in range(3):
        f_all = eng.eval(synthetic.perturbed_candidates(x0, B, first=100 * call))
        for r in range(world):
            if call == 2:    # the last rank sent nothing: its slice still holds call 0's values (same half of the array)
                from optimo_b200.core import shard_range
                b0, nb = shard_range(B, world - 1, world)
                assert np.array_equal(got[r][call][:b0], f_all[:b0])
                assert np.array_equal(got[r][call][b0:], got[r][0][b0:])
            else:
                assert np.array_equal(got[r][call], f_all), (call, r)


def test_peer_gather_wait_times_out_when_a_rank_never_delivers(tmp_path):
    """A rank that maps its peers and then never evaluates (a crashed process, seen from the others) must not hang the GPU: the
    waiting kernel gives up after ~2 s, the call returns OMO_NCCL_ERROR, and the context keeps working."""
    import subprocess
    import sys
    n = _n_gpus()
    B, world = 64, 2
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "peer_gather_worker.py")
    procs = [subprocess.Popen([sys.executable, worker, str(tmp_path), str(r), str(world), str(r if n >= world else 0), str(B), "absent"],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(world)]
    logs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=180)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        logs.append(out.decode(errors="replace")[-2000:])
    assert all(p.returncode == 0 for p in procs), logs
    with open(os.path.join(str(tmp_path), "gave_up0")) as fh:
        outcome = fh.read()
    assert outcome.startswith("error after") and "status 3" in outcome and "did not arrive" in outcome, outcome
    assert float(outcome.split()[2]) < 30.0, outcome
    c = Core(synthetic.stick_240(load_scene("stick-figure")), device=0)
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    eng.set_batch_hint(B)
    from optimo_b200.core import shard_range
    b0, nb = shard_range(B, 0, world)
    assert np.array_equal(np.load(os.path.join(str(tmp_path), "after0.npy")), eng.eval(synthetic.perturbed_candidates(x0, B)[b0:b0 + nb]))

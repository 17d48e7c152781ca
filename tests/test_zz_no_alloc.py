"""omo_eval_device_static (include/omo.h): the batch objective with its scratch in the library's file-scope `static __device__`
array -- for callers that cannot call cudaMalloc. It has no counterpart in the reference; it is checked against the existing
entry point omo_eval_device through the C ABI, on two calls in a row (the array keeps the first call's contents).

Collected last on purpose (a failure here must not hide the other tests); passes alone in a fresh process
and after the rest of the suite."""
import numpy as np
import pytest

from conftest import load_scene, scene_path

from optimo_b200 import synthetic
from optimo_b200.core import Core, EngineError

pytestmark = pytest.mark.gpu

# the two entry points run the same device code; where omo_eval_device takes the scene-specialised kernel (launches of >= 64 K
# sample-evals) and the static one the table-driven kernel they differ in the last ulps (include/omo.h: ~1e-13)
RTOL = 1e-12


def _both(eng, X, want_terms=True):
    """f (and terms) of the rows of X through omo_eval_device and through omo_eval_device_static, the caller's buffers distinct."""
    B, D = X.shape
    dX = eng.device_alloc(8 * B * D)
    eng.h2d(dX, X)
    out = []
    for call in (eng.eval_device, eng.eval_device_static):
        dF, dT = eng.device_alloc(8 * B), eng.device_alloc(24 * B)
        jit0 = eng.jit_info()["launches"]
        call(dX, B, dF, dT if want_terms else None, sync=True)
        out.append((eng.d2h(dF, (B,)), eng.d2h(dT, (B, 3)) if want_terms else None, eng.jit_info()["launches"] > jit0))
        eng.device_free(dF); eng.device_free(dT)
    eng.device_free(dX)
    assert not out[1][2], "the static entry point must stay on the table-driven kernel (a module load allocates)"
    return out


def _scene(name):
    if name == "stick240":
        return Core(synthetic.stick_240(load_scene("stick-figure"))), 12.5
    return Core(scene_path(name)), 12.5


@pytest.mark.parametrize("name", ["stick240", "conductor"])
def test_static_scratch_entry_point_matches_omo_eval_device_on_two_calls_in_a_row(name):
    c, w_traj = _scene(name)
    c.UpdateGlobalWeight(w_traj)                        # D_traj on: the FK steps and p_original are part of the launch
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))              # the allocating set-up: derived state built by one ordinary evaluation
    big = 220 if name == "stick240" else 64             # stick240: 220 x 1216 sample-evals -> omo_eval_device takes the specialised kernel
    calls = [synthetic.perturbed_candidates(x0, big), synthetic.perturbed_candidates(x0, 5, first=1000)]
    launches0 = eng.counters()["launches"]
    for i, X in enumerate(calls):                       # second call: fewer rows over the first call's leftovers
        (f_ref, t_ref, ref_specialised), (f_st, t_st, _) = _both(eng, X)
        assert np.all(np.isfinite(f_st)) and np.all(f_st > 0.0)
        assert np.allclose(f_st, f_ref, rtol=RTOL, atol=0.0), (i, float(np.max(np.abs(f_st / f_ref - 1.0))))
        assert np.allclose(t_st, t_ref, rtol=RTOL, atol=RTOL * float(np.max(np.abs(t_ref))))
        if not ref_specialised:
            assert np.array_equal(f_st, f_ref) and np.array_equal(t_st, t_ref)   # both on the table-driven kernel: the same bits
    assert eng.counters()["launches"] > launches0
    # the same rows again, without terms: identical to the call before (nothing depends on what the array held)
    f_again = _both(eng, calls[1], want_terms=False)[1][0]
    f_first = _both(eng, calls[1], want_terms=False)[1][0]
    assert np.array_equal(f_again, f_first)


def test_static_scratch_entry_point_allocates_no_device_memory():
    torch = pytest.importorskip("torch")
    c, w = _scene("stick240")
    x0 = c.GetDescriptor()
    lo, hi = c.window
    eng = c.engine(synced_window=(lo, hi))
    B = 100
    X = synthetic.perturbed_candidates(x0, B)
    dX, dF = eng.device_alloc(8 * X.size), eng.device_alloc(8 * B)
    eng.h2d(dX, X)
    eng.eval_device_static(dX, B, dF, None, sync=True)  # first launch of a kernel may load it lazily
    free0 = torch.cuda.mem_get_info(0)[0]
    for _ in range(3):
        eng.eval_device_static(dX, B, dF, None, sync=True)
    assert torch.cuda.mem_get_info(0)[0] == free0
    eng.device_free(dX); eng.device_free(dF)


def test_static_scratch_entry_point_rejects_what_does_not_fit_and_what_is_not_set_up():
    c, _ = _scene("stick240")
    x0 = c.GetDescriptor()
    lo, hi = c.window
    raw = c.engine()                                    # no window / weights / x_original pushed yet
    dX, dF = raw.device_alloc(8 * 4096 * x0.size), raw.device_alloc(8 * 4096)
    with pytest.raises(EngineError, match="status 5"):  # OMO_NOT_READY
        raw.eval_device_static(dX, 4, dF, None, sync=True)
    eng = c.engine(synced_window=(lo, hi))
    with pytest.raises(EngineError, match=r"status 1: .*static array holds 8388608 \(at most \d+ candidates"):   # OMO_BAD_ARG
        eng.eval_device_static(dX, 4096, dF, None, sync=True)
    X = synthetic.perturbed_candidates(x0, 4)
    eng.h2d(dX, X)
    eng.eval_device_static(dX, 4, dF, None, sync=True)  # the context is still usable after the refusals
    assert np.allclose(eng.d2h(dF, (4,)), eng.eval(X), rtol=RTOL, atol=0.0)
    eng.device_free(dX); eng.device_free(dF)

"""bp_tail_static (include/bp_b200.h): the fused task tail for callers that may not allocate - scratch in a file-scope
device array, no cudaMalloc.  No counterpart in the reference, so the
check is against the library's own allocating entry point, bp_get_values_tail, bit for bit, on calls in a row that
share the static scratch.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

from conftest import split_cam

MAX_VALUES = 1 << 27  # BP_TAIL_STATIC_MAX_VALUES


def test_tail_static_rejects_volumes_beyond_the_static_scratch():
    """Argument checks run before any CUDA call: no GPU needed."""
    from plant3dvision_b200 import _cabi
    L = _cabi.lib()
    codes = (ctypes.c_int32 * 3)(-1, 1, 0)
    dummy = ctypes.c_void_p(256)
    rc = L.bp_tail_static(0, _cabi.MODE_CARVING, dummy, (MAX_VALUES // 1024) + 1, 1024, 1024, codes, 0, dummy, dummy, None)
    assert rc == -1  # BP_EINVAL
    msg = L.bp_last_error().decode() if isinstance(L.bp_last_error(), bytes) else str(L.bp_last_error())
    assert "limit" in msg and str(MAX_VALUES) in msg, msg
    assert L.bp_tail_static(0, _cabi.MODE_CARVING, None, 4, 4, 4, codes, 0, dummy, dummy, None) == -1
    assert L.bp_tail_static(0, _cabi.MODE_CARVING, dummy, 4, 4, 4, codes, _cabi.TAIL_EXP_CLIP, dummy, dummy, None) == -1
    assert L.bp_tail_static(0, _cabi.MODE_AVERAGING, dummy, 4, 8, 4, None, 0, dummy, dummy, None) == -1  # pitch < nz


def _scene(scale, n_views):
    from plant3dvision_b200 import synth
    w = synth.workload("C5", scale)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    return w, [(*split_cam(w["cams"][v]), masks[v]) for v in range(n_views)]


def _static_tail(bp, float64, exp_clip, codes):
    """The volume and (min, max) through bp_tail_static, on torch-owned device buffers (the caller's outputs)."""
    import torch
    from plant3dvision_b200 import _cabi
    ptr, _, pitch = bp.device_state()
    ni, ny, nz = (int(x) for x in bp.local_shape)
    carving = bp._mode == _cabi.MODE_CARVING
    dt = torch.float64 if float64 else (torch.int32 if carving else torch.float32)
    out = torch.full((ni * ny * nz,), 77, dtype=dt, device="cuda:0")
    stats = torch.full((2,), 12345.0, dtype=torch.float64, device="cuda:0")
    bp.sync()                 # the handle's own stream has finished the views
    torch.cuda.synchronize()  # and the fills above are done: the call below runs on the default stream
    flags = (_cabi.TAIL_F64 if float64 else 0) | (_cabi.TAIL_EXP_CLIP if exp_clip else 0)
    c = (ctypes.c_int32 * 3)(*codes) if codes is not None else None
    _cabi.check(_cabi.lib().bp_tail_static(0, bp._mode, ctypes.c_void_p(ptr), ni * ny, nz, pitch, c, flags,
                                           ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(stats.data_ptr()), None))
    torch.cuda.synchronize()
    lo, hi = stats.cpu().tolist()
    return out.cpu().numpy().reshape(ni, ny, nz), lo, hi


@pytest.mark.gpu
def test_tail_static_equals_get_values_tail_on_calls_in_a_row():
    from plant3dvision_b200.cl import Backprojection
    w, views = _scene(0.1, 24)
    carve = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving", default_value=0)
    carve.process_views(views)
    w2, views2 = _scene(0.06, 12)  # a smaller volume next: fewer partials than the scratch still holds from before
    avg = Backprojection(w2["shape"], w2["origin"], w2["voxel_size"], type="averaging", log=True)
    avg.process_views(views2)
    cases = [(carve, False, False, (-1, 1, 0)), (avg, False, True, None), (carve, True, False, (-1, 1, 0)),
             (avg, True, False, None), (avg, False, False, None), (carve, False, False, (-1, 1, 0))]
    for bp, f64, ec, codes in cases:
        want, lo, hi = bp.get_values_tail(float64=f64, exp_clip=ec)
        want = np.array(want, copy=True)
        for attempt in range(2):  # two calls in a row on the same input: the shared scratch keeps no state that matters
            got, glo, ghi = _static_tail(bp, f64, ec, codes)
            assert got.dtype == want.dtype
            assert np.array_equal(got, want), f"{(got != want).sum()} values differ (f64={f64} exp_clip={ec} call {attempt})"
            assert (glo, ghi) == (lo, hi), ((glo, ghi), (lo, hi))
    assert lo == -1.0 and hi == 1.0  # the carved volume holds carved and seen voxels
    carve.close()
    avg.close()

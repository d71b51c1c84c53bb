"""Speculative read-back of the carve volume (bp_set_readback_target; default since round 2, ``BP_SPEC_READBACK=0`` or
``Backprojection(..., speculative_readback=False)`` switches it off).  With host masks arriving in several staging
chunks the library sends an early copy of the volume to the pinned host buffer while the remaining masks are still
being uploaded, and get_values() only writes the rows that changed since.  The values must be exactly what
get_values() returns without it.  First run on a B200 in round 2 (profiles/r02_pytest_spec_readback.log).
"""
import os

import numpy as np
import pytest

from conftest import split_cam

pytestmark = pytest.mark.gpu


def _scene(scale=0.064, step=4):
    from plant3dvision_b200 import synth
    w = synth.workload("C5", scale)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    views = [(*split_cam(w["cams"][v]), masks[v]) for v in range(0, len(w["cams"]), step)]
    return w, views


def _oracle(w, views, default_value=0):
    from oracle.backprojection_oracle import BackprojectionOracle
    ora = BackprojectionOracle(w["shape"], w["origin"], w["voxel_size"], type="carving", default_value=default_value)
    for v in views:
        ora.process_view(*v)
    return ora.get_values()


@pytest.fixture()
def spec_on(monkeypatch):
    monkeypatch.setenv("BP_SPEC_READBACK", "1")


@pytest.mark.parametrize("default_value", [0, 5, -1])
@pytest.mark.parametrize("at", [None, "0.1", "0.5"])
def test_speculative_readback_equals_plain(spec_on, monkeypatch, default_value, at):
    from plant3dvision_b200.cl import Backprojection
    w, views = _scene()
    monkeypatch.setenv("BP_STAGE_BYTES", str(3 * views[0][3].nbytes))  # 3 views per chunk -> 20 chunks
    if at is not None:
        monkeypatch.setenv("BP_SPEC_AT", at)
    want = _oracle(w, views, default_value)
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving", default_value=default_value)
    for rep in range(3):  # the second and third rounds re-use the snapshot buffers
        if rep:
            bp.clear()
        bp.process_views(views)
        got = bp.get_values()
        assert np.array_equal(got, want), f"round {rep}: {(got != want).sum()} voxels differ"
        patched, total = bp.spec_stats()
        assert total == w["shape"][0] * w["shape"][1]
        assert 0 <= patched < total  # an early copy was taken, and only some rows had to be re-sent
    bp.close()


def test_other_readers_and_orders(spec_on, monkeypatch):
    from plant3dvision_b200.cl import Backprojection
    w, views = _scene()
    monkeypatch.setenv("BP_STAGE_BYTES", str(3 * views[0][3].nbytes))
    want = _oracle(w, views)
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving")
    # the tail reader while an early copy is pending, then the plain one
    bp.process_views(views)
    tail, lo, hi = bp.get_values_tail(float64=True)
    assert np.array_equal(tail, want) and lo == want.min() and hi == want.max()
    assert np.array_equal(bp.get_values(), want)
    # pending copy, then more views in ONE chunk (no new early copy), then the read-back
    bp.clear()
    bp.process_views(views[:-2])
    monkeypatch.setenv("BP_STAGE_BYTES", str(1 << 30))
    bp.process_views(views[-2:])
    assert np.array_equal(bp.get_values(), want)
    # pending copy, clear, read-back of the seeded grid
    monkeypatch.setenv("BP_STAGE_BYTES", str(3 * views[0][3].nbytes))
    bp.process_views(views)
    bp.clear()
    assert (bp.get_values() == 0).all()
    # per-view calls and device-resident masks never speculate
    for v in views:
        bp.process_view(*v)
    assert np.array_equal(bp.get_values(), want)
    bp.close()


def test_target_must_be_pinned():
    import ctypes
    from plant3dvision_b200 import _cabi
    from plant3dvision_b200.cl import Backprojection
    bp = Backprojection([8, 8, 8], [0, 0, 0], 1.0, type="carving")
    pageable = np.zeros(512, dtype=np.int32)
    rc = _cabi.lib().bp_set_readback_target(bp._handle(), ctypes.c_void_p(pageable.ctypes.data))
    assert rc == -1 and b"page-locked" in _cabi.lib().bp_last_error()
    avg = Backprojection([8, 8, 8], [0, 0, 0], 1.0, type="averaging")
    rc = _cabi.lib().bp_set_readback_target(avg._handle(), ctypes.c_void_p(avg.values_h.ctypes.data))
    assert rc == -4
    bp.close()
    avg.close()

"""The tail of ``Voxels.run`` fused into the read-back: ``bp_get_values_tail`` against numpy doing
what tasks/cl.py:166-174 and cl.py:249-254 do on the host.

Bars: the un-transformed values, the float64 conversion and min / max are exact.  ``exp`` is the one transcendental
of the path: numpy's and CUDA's are both within 1 ulp of the true value, so they are gated at 4 ulp of the output
dtype (float64: 9e-16 relative; float32: 4.8e-7), far inside north_star's 1e-5; the clip to 1 must be exact.
"""
import json

import numpy as np
import pytest

from conftest import CHANNELS, split_cam, vp_shape_origin

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def BP():
    from plant3dvision_b200.cl import Backprojection
    return Backprojection


def views_of(vp_inputs, channel, invert=False):
    maps = vp_inputs["maps"][channel]
    return [(*split_cam(vp_inputs["cams"][v]), np.invert(maps[v]) if invert else maps[v]) for v in range(len(maps))]


@pytest.mark.parametrize("log", [True, False])
def test_averaging_tail_equals_numpy_tail(BP, vp_inputs, log):
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.6)
    bp = BP(shape, origin, 0.6, type="averaging", log=log)
    bp.process_views(views_of(vp_inputs, "stem"))
    raw = bp.get_values().copy()
    v32, lo, hi = bp.get_values_tail()
    assert v32.dtype == np.float32 and np.array_equal(v32, raw)
    assert lo == float(raw.min()) and hi == float(raw.max())
    v64, lo2, hi2 = bp.get_values_tail(float64=True)
    assert v64.dtype == np.float64 and np.array_equal(v64, raw.astype(np.float64)) and (lo2, hi2) == (lo, hi)
    # float64 stack of process_fileset, then np.exp and the clip (tasks/cl.py:172-174)
    want = np.exp(raw.astype(np.float64))
    want[want > 1] = 1.0
    out = np.empty((2, *shape))
    got, lo3, hi3 = bp.get_values_tail(out=out[1], float64=True, exp_clip=True)
    assert np.shares_memory(got, out) and (lo3, hi3) == (lo, hi)  # min / max are those of the RAW volume
    assert got.max() <= 1.0 and np.array_equal(got == 1.0, want == 1.0)
    np.testing.assert_allclose(got, want, rtol=4 * np.finfo(np.float64).eps, atol=0)
    # no labels: the volume stays float32 and np.exp runs in float32
    want32 = np.exp(raw)
    want32[want32 > 1] = 1.0
    got32, _, _ = bp.get_values_tail(exp_clip=True)
    assert got32.dtype == np.float32 and got32.max() <= 1.0
    np.testing.assert_allclose(got32, want32, rtol=4 * np.finfo(np.float32).eps, atol=0)
    with pytest.raises(ValueError):
        bp.get_values_tail(out=np.empty(shape, np.float32), float64=True)
    bp.close()


def test_carving_tail(BP, vp_inputs):
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.6)
    bp = BP(shape, origin, 0.6, type="carving")
    v0, lo, hi = bp.get_values_tail()
    assert (lo, hi) == (0.0, 0.0) and not v0.any()  # the untouched grid is what tasks/cl.py:168 warns about
    bp.process_views(views_of(vp_inputs, "background", invert=True))
    raw = bp.get_values().copy()
    v, lo, hi = bp.get_values_tail()
    assert v.dtype == np.int32 and np.array_equal(v, raw) and (lo, hi) == (raw.min(), raw.max()) == (-1, 1)
    v64, _, _ = bp.get_values_tail(float64=True)
    assert v64.dtype == np.float64 and np.array_equal(v64, raw)
    with pytest.raises(ValueError, match="averaging"):
        bp.get_values_tail(exp_clip=True)
    bp.close()


def test_nan_propagates_into_min_max_like_numpy(BP):
    bp = BP([8, 9, 10], [0, 0, 0], 1.0, type="averaging")
    cam = (np.array([10, 10, 8, 8], np.float32), np.eye(3, dtype=np.float32).ravel(), np.array([0, 0, 5], np.float32))
    m = np.full((16, 16), np.nan, dtype=np.float32)
    bp.process_view(*cam, m)
    raw = bp.get_values().copy()
    assert np.isnan(raw).any()
    _, lo, hi = bp.get_values_tail()
    assert np.isnan(lo) and np.isnan(hi) and np.isnan(raw.min())
    bp.close()


def _write_scan(tmp_path, vp_inputs, channels, n_views=18):
    from PIL import Image
    scan_dir = tmp_path / "scan"
    (scan_dir / "Segmentation2D_1").mkdir(parents=True)
    (scan_dir / "metadata" / "Segmentation2D_1").mkdir(parents=True)
    files = []
    for v in range(n_views):
        c = vp_inputs["cams"][v]
        cam = {"camera_model": {"params": [float(x) for x in c[0:4]] + [0, 0, 0, 0]},
               "rotmat": c[4:13].reshape(3, 3).tolist(), "tvec": c[13:16].tolist()}
        for ch in channels:
            fid = f"{v:05d}_{ch}"
            Image.fromarray(vp_inputs["maps"][ch][v]).save(scan_dir / "Segmentation2D_1" / f"{fid}.png")
            with open(scan_dir / "metadata" / "Segmentation2D_1" / f"{fid}.json", "w") as f:
                json.dump({"channel": ch, "camera": cam}, f)
            files.append({"id": fid, "file": f"{fid}.png"})
    with open(scan_dir / "metadata" / "Segmentation2D_1.json", "w") as f:
        json.dump({"label_names": list(channels)}, f)
    with open(scan_dir / "metadata" / "metadata.json", "w") as f:
        json.dump({"bounding_box": vp_inputs["bbox"]}, f)
    with open(scan_dir / "files.json", "w") as f:
        json.dump({"filesets": [{"id": "Segmentation2D_1", "files": files}]}, f)
    return scan_dir


def test_voxels_task_writes_what_the_reference_tail_writes(tmp_path, vp_inputs):
    """ml_pipe_virtual.toml [Voxels]: averaging, log (task default), labels from the fileset metadata, NPZ output
    (tasks/cl.py:149-186) - against the oracle followed by numpy's exp / clip."""
    from oracle.backprojection_oracle import BackprojectionOracle
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200.tasks import cl as tcl
    channels = ["stem", "leaf", "background"]
    scan_dir = _write_scan(tmp_path, vp_inputs, channels, n_views=6)
    out = tcl.run_voxels(str(scan_dir), {"type": "averaging", "voxel_size": 0.6, "log": True, "camera_metadata": "camera",
                                         "upstream_mask": "Segmentation2D"})
    assert out.filename == "Voxels.npz" and out.get_metadata("voxel_size") == 0.6
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.6)
    assert out.get_metadata("origin") == [float(o) for o in origin]
    fs = fsdb_lite.Scan(str(scan_dir)).get_fileset("Segmentation2D_1")
    ora = BackprojectionOracle(shape, origin, 0.6, type="averaging", labels=channels, log=True)
    want = np.exp(ora.process_fileset(fs.get_files(), "camera"))
    want[want > 1] = 1.0
    with np.load(out.path()) as z:
        assert z.files == channels
        for i, ch in enumerate(channels):
            assert z[ch].dtype == np.float64 and z[ch].shape == tuple(shape)
            np.testing.assert_allclose(z[ch], want[i], rtol=4 * np.finfo(np.float64).eps, atol=0)
            assert np.array_equal(z[ch] == 1.0, want[i] == 1.0)

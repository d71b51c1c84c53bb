"""View rectangles (default since round 2; ``BP_VIEW_RECT=0`` switches them off): the library copies and packs only the
rectangle of every view that the handle's planes project into (bp_view_rect; conservativeness is pinned on the CPU by
tests/test_view_rect.py).  Results must not change at all.  First run on a B200 in round 2
(profiles/r02_pytest_view_rect.log), then made the default: C5 uploads 3.54 GB instead of 5.76 GB per step.
"""
import os

import numpy as np
import pytest

from conftest import split_cam

pytestmark = pytest.mark.gpu


@pytest.fixture()
def rect_on(monkeypatch):
    monkeypatch.setenv("BP_VIEW_RECT", "1")


def _workload(name="C5", scale=0.064, soft=0.0):
    from plant3dvision_b200 import synth
    w = synth.workload(name, scale)
    assert w["W"] % 16 == 0  # the batched uint8 packer (the one that knows rectangles) needs 16-byte rows
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"], soft_px=soft).numpy()
    return w, [(*split_cam(w["cams"][v]), masks[v]) for v in range(len(w["cams"]))]


def _oracle(w, views, kind, log=False):
    from oracle.backprojection_oracle import BackprojectionOracle
    ora = BackprojectionOracle(w["shape"], w["origin"], w["voxel_size"], type=kind, log=log)
    for v in views:
        ora.process_view(*v)
    return ora.get_values()


@pytest.mark.parametrize("kind", ["carving", "averaging"])
def test_host_masks_trimmed_upload_is_bit_identical(rect_on, kind):
    from plant3dvision_b200.cl import Backprojection
    w, views = _workload(soft=2.0 if kind == "averaging" else 0.0)
    views = views[::4]
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type=kind, log=(kind == "averaging"))
    # poison the staging buffers and the tile pool first: whatever is left outside the rectangles must never be read
    poison = [(i, r, t, np.zeros_like(m) if kind == "carving" else np.full_like(m, 255)) for (i, r, t, m) in views]
    bp.process_views(poison)
    bp.clear()
    before = bp.ingest_bytes()
    bp.process_views(views)
    got = bp.get_values().copy()
    moved = bp.ingest_bytes() - before
    bp.close()
    want = _oracle(w, views, kind, log=(kind == "averaging"))
    assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ"
    full = sum(m.nbytes for (_, _, _, m) in views)
    assert moved < 0.85 * full, (moved, full)


def test_float_maps_and_streaming_views(rect_on):
    from plant3dvision_b200.cl import Backprojection
    w, views = _workload(soft=2.0)
    views = [(i, r, t, (m.astype(np.float32) / 255.0)) for (i, r, t, m) in views[::6]]
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="averaging", log=False)
    for v in views:
        bp.process_view(*v)
    got = bp.get_values().copy()
    bp.close()
    assert np.array_equal(got, _oracle(w, views, "averaging"))


def test_device_masks(rect_on):
    import torch
    from plant3dvision_b200.cl import Backprojection
    w, views = _workload()
    views = views[::4]
    dev = [torch.from_numpy(m).cuda() for (_, _, _, m) in views]
    cams = np.stack([np.concatenate([i, r, t]) for (i, r, t, _) in views]).astype(np.float32)
    hw = np.array([[w["H"], w["W"]]] * len(views), dtype=np.int32)
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving")
    bp.process_device_views(cams, [d.data_ptr() for d in dev], hw)
    got = bp.get_values().copy()
    assert bp.ingest_bytes() == 0
    bp.close()
    assert np.array_equal(got, _oracle(w, views, "carving"))


def test_slabs_and_stripes(rect_on):
    from plant3dvision_b200.cl import Backprojection
    w, views = _workload()
    views = views[::4]
    shape = w["shape"]
    want = _oracle(w, views, "carving")
    for a, b in ((0, 5), (5, shape[0] // 2), (shape[0] // 2, shape[0])):
        s = Backprojection(shape, w["origin"], w["voxel_size"], type="carving", i_range=(a, b))
        s.process_views(views)
        np.testing.assert_array_equal(s.get_values(), want[a:b])
        s.close()
    for rank in range(3):
        s = Backprojection(shape, w["origin"], w["voxel_size"], type="carving", stripes=(2, 3, rank))
        s.process_views(views)
        np.testing.assert_array_equal(s.get_values(), want[s.planes])
        s.close()


def test_camera_inside_the_grid_and_mixed_sizes(rect_on):
    from plant3dvision_b200.cl import Backprojection
    from oracle.backprojection_oracle import BackprojectionOracle
    rng = np.random.default_rng(3)
    shape, origin, vs = [40, 36, 70], [-20.0, -18.0, -35.0], 1.0
    views = []
    for n, (H, W) in enumerate([(96, 160), (130, 208), (64, 64), (200, 320)]):
        ang = 2 * np.pi * n / 4
        dist = [120.0, 60.0, 5.0, 200.0][n]  # the third camera sits inside the grid: whole-image fallback
        pos = np.array([dist * np.cos(ang), dist * np.sin(ang), 3.0])
        zc = -pos / np.linalg.norm(pos)
        xc = np.cross([0, 0, 1.0], zc)
        xc /= np.linalg.norm(xc)
        R = np.stack([xc, np.cross(zc, xc), zc])
        t = -R @ pos
        f = 0.8 * W
        m = (rng.random((H, W)) < 0.9).astype(np.uint8) * 255
        views.append((np.array([f, f, W / 2, H / 2], np.float32), R.reshape(-1).astype(np.float32), t.astype(np.float32), m))
    bp = Backprojection(shape, origin, vs, type="carving")
    bp.process_views(views)
    got = bp.get_values().copy()
    bp.close()
    ora = BackprojectionOracle(shape, origin, vs, type="carving")
    for v in views:
        ora.process_view(*v)
    assert np.array_equal(got, ora.get_values())


@pytest.mark.parametrize("kind,sharding", [("carving", "slabs"), ("carving", "stripes"), ("averaging", "stripes")])
def test_two_gpus_sharded_ingest(rect_on, tmp_path, kind, sharding):
    """The sharded ingest of parallel.py with rectangles: a rank packs its share of the views trimmed to the WHOLE
    grid's rectangle (bp_pack_views_cams), the tiles are all-gathered, every rank carves its planes from them."""
    import socket
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    from _gloo_worker import scene, worker
    from oracle.backprojection_oracle import BackprojectionOracle
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(worker, args=(2, port, kind, sharding, 0, str(tmp_path), "nccl"), nprocs=2, join=True)  # inherits the env
    shape, origin, vs, views = scene()
    o = BackprojectionOracle(shape, origin, vs, type=kind)
    for v in views:
        o.process_view(*v)
    for rank in range(2):
        assert np.array_equal(np.load(tmp_path / f"rank{rank}.npz")["vol"], o.get_values())

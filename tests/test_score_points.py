"""SegmentedPointCloud scoring: ``plant3dvision_b200.proc3d.score_points`` (one CUDA launch) against
the oracle (``oracle/vol2pcd_oracle.score_points`` = tasks/proc3d.py:203-230 with proc3d.py:655-659).

Bar: scores are sums of uint8 map values, i.e. exact integers, and labels are their arg-max: both must be EQUAL.
The only unpinned arithmetic is the reference's ``rot @ pts.T`` (BLAS, may fuse multiply-adds) against the device's
separate float64 multiplies and adds; that can move a projected coordinate by an ulp, which matters only if it
sits within an ulp of a pixel boundary (probability ~1e-13 per projection; the seeded cases below are not there).
"""
import json
import os

import numpy as np
import pytest

from conftest import CHANNELS

LABELS = [c for c in CHANNELS if c != "background"]


def camera_dict(c):
    return {"rotmat": np.asarray(c[4:13]).reshape(3, 3).tolist(), "tvec": np.asarray(c[13:16]).tolist(),
            "camera_model": {"params": [float(c[0]), float(c[1]), float(c[2]), float(c[3])]}}


def vp_views(vp_inputs, channels=CHANNELS, n_views=None):
    """File order of a Segmentation2D fileset: all channels of image 0, then image 1, ..."""
    cams = vp_inputs["cams"]
    n = len(cams) if n_views is None else n_views
    return [(ch, camera_dict(cams[v]), vp_inputs["maps"][ch][v]) for v in range(n) for ch in channels]


def vp_points(vp_inputs, n, seed=0, spread=1.3):
    rng = np.random.default_rng(seed)
    b = vp_inputs["bbox"]
    lo = np.array([b["x"][0], b["y"][0], b["z"][0]])
    hi = np.array([b["x"][1], b["y"][1], b["z"][1]])
    mid, half = (lo + hi) / 2, (hi - lo) / 2
    return mid + (rng.random((n, 3)) * 2 - 1) * half * spread


def reference_loop(points, views, labels):
    """tasks/proc3d.py:203-230 with its per-point Python loop kept as written (small cases only)."""
    from oracle.vol2pcd_oracle import backproject_points
    pts = np.asarray(points)
    scores = np.zeros((len(labels), len(pts)))

    def is_in_pict(px, shape):
        return px[0] >= 0 and px[0] < shape[1] and px[1] >= 0 and px[1] < shape[0]

    for label, camera, mask in views:
        if label not in labels:
            continue
        if camera is None:
            continue
        rotmat = np.array(camera["rotmat"])
        tvec = np.array(camera["tvec"])
        intrinsics = camera["camera_model"]["params"]
        K = np.array([[intrinsics[0], 0, intrinsics[2]], [0, intrinsics[1], intrinsics[3]], [0, 0, 1]])
        pixels = np.asarray(backproject_points(pts, K, rotmat, tvec) + 0.5, dtype=int)
        label_idx = labels.index(label)
        for i, px in enumerate(pixels):
            if is_in_pict(px, mask.shape):
                scores[label_idx, i] += mask[px[1], px[0]]
    return np.argmax(scores, axis=0).flatten(), scores


# ---------------------------------------------------------------------------------------------
# CPU
# ---------------------------------------------------------------------------------------------
def test_oracle_vectorised_form_equals_the_reference_loop(vp_inputs):
    from oracle.vol2pcd_oracle import score_points
    pts = vp_points(vp_inputs, 300, seed=1, spread=2.5)  # some points project outside the pictures
    views = vp_views(vp_inputs, n_views=4)
    views[3] = (views[3][0], None, views[3][2])  # a view without a pose is skipped
    l0, s0 = reference_loop(pts, views, LABELS)
    l1, s1 = score_points(pts, views, LABELS)
    assert np.array_equal(s0, s1) and np.array_equal(l0, l1)
    assert s0.max() > 0 and (s0 == 0).any()


def test_backproject_points_is_the_pinhole_projection(vp_inputs):
    from oracle.vol2pcd_oracle import backproject_points as o_bp
    import importlib
    bp = importlib.import_module("plant-3d-vision_b200.proc3d").backproject_points
    c = vp_inputs["cams"][2]
    K = np.array([[c[0], 0, c[2]], [0, c[1], c[3]], [0, 0, 1]])
    R, t = c[4:13].reshape(3, 3), c[13:16]
    pts = vp_points(vp_inputs, 50)
    uv = o_bp(pts, K, R, t)
    assert np.array_equal(uv, bp(pts, K, R, t))
    xc = pts @ R.T + t
    assert np.allclose(uv[:, 0], c[0] * xc[:, 0] / xc[:, 2] + c[2], rtol=1e-12)
    assert np.allclose(uv[:, 1], c[1] * xc[:, 1] / xc[:, 2] + c[3], rtol=1e-12)


# ---------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------
def _proc3d():
    import importlib
    return importlib.import_module("plant-3d-vision_b200.proc3d")


@pytest.mark.gpu
@pytest.mark.parametrize("n,spread", [(1, 1.0), (257, 1.0), (20000, 1.3), (5000, 4.0)])
def test_score_points_equals_oracle(vp_inputs, n, spread):
    from oracle.vol2pcd_oracle import score_points as oracle_score
    pts = vp_points(vp_inputs, n, seed=n, spread=spread)
    views = vp_views(vp_inputs)
    want_l, want_s = oracle_score(pts, views, LABELS)
    got_l, got_s, ms = _proc3d().score_points(pts, views, LABELS)
    assert got_s.shape == want_s.shape and got_s.dtype == np.float64
    assert np.array_equal(got_s, want_s)
    assert np.array_equal(got_l, want_l)
    assert ms > 0


@pytest.mark.gpu
def test_score_points_edge_cases(vp_inputs):
    from oracle.vol2pcd_oracle import score_points as oracle_score
    p3 = _proc3d()
    views = vp_views(vp_inputs, n_views=5)
    # points on the optical centre plane (z_c = 0), behind the cameras, non-finite, and far away
    c = vp_inputs["cams"][0]
    R, t = c[4:13].reshape(3, 3), c[13:16]
    centre = -R.T @ t
    pts = np.concatenate([vp_points(vp_inputs, 64, seed=3), centre[None], (centre - 50 * R[2])[None],
                          [[np.nan, 0, 0], [np.inf, 1, 2], [1e300, -1e300, 1e300], [0, 0, 0]]])
    views[4] = (views[4][0], None, views[4][2])                       # no pose: skipped
    views.append(("rgb", views[0][1], views[0][2]))                  # label not in the list: skipped
    want_l, want_s = oracle_score(pts, views, LABELS)
    got_l, got_s, _ = p3.score_points(pts, views, LABELS)
    assert np.array_equal(got_s, want_s) and np.array_equal(got_l, want_l)
    # label order decides ties (np.argmax: first maximum): all-zero scores -> label 0
    rev = LABELS[::-1]
    got_l2, got_s2, _ = p3.score_points(pts, views, rev)
    assert np.array_equal(got_s2, got_s[::-1])
    assert (got_l2[(got_s == 0).all(axis=0)] == 0).all()
    # empty inputs
    l0, s0, _ = p3.score_points(np.zeros((0, 3)), views, LABELS)
    assert l0.shape == (0,) and s0.shape == (len(LABELS), 0)
    l1, s1, _ = p3.score_points(pts, [], LABELS)
    assert (l1 == 0).all() and (s1 == 0).all()
    # maps of different sizes in one call
    small = [(l, cam, m[:100, :333].copy()) if i % 2 else (l, cam, m) for i, (l, cam, m) in enumerate(views)]
    want_l, want_s = oracle_score(pts, small, LABELS)
    got_l, got_s, _ = p3.score_points(pts, small, LABELS)
    assert np.array_equal(got_s, want_s) and np.array_equal(got_l, want_l)
    with pytest.raises(TypeError):
        p3.score_points(pts, [(LABELS[0], views[0][1], np.zeros((4, 4), np.float32))], LABELS)


@pytest.mark.gpu
def test_segmented_point_cloud_task_on_a_scan_directory(vp_inputs, tmp_path):
    """The task mirror end to end: PointCloud.ply + Segmentation2D PNGs with 'channel' / 'camera' metadata in an
    FSDB directory -> SegmentedPointCloud.ply with colours and the 'labels' metadata (tasks/proc3d.py:185-253)."""
    import importlib
    from PIL import Image
    from oracle.vol2pcd_oracle import score_points as oracle_score
    tasks = importlib.import_module("plant-3d-vision_b200.tasks.proc3d")
    scan_dir = tmp_path / "scan"
    (scan_dir / "PointCloud_1").mkdir(parents=True)
    (scan_dir / "Segmentation2D_1").mkdir()
    (scan_dir / "metadata" / "Segmentation2D_1").mkdir(parents=True)
    pts = vp_points(vp_inputs, 3000, seed=9, spread=1.0)
    nrm = np.tile([0.0, 0.0, 1.0], (len(pts), 1))
    tasks.write_ply(scan_dir / "PointCloud_1" / "PointCloud.ply", pts, nrm)
    files, views = [], []
    for v in range(6):
        for ch in CHANNELS:
            fid = f"{v:05d}_{ch}"
            Image.fromarray(vp_inputs["maps"][ch][v]).save(scan_dir / "Segmentation2D_1" / f"{fid}.png")
            cam = camera_dict(vp_inputs["cams"][v])
            with open(scan_dir / "metadata" / "Segmentation2D_1" / f"{fid}.json", "w") as f:
                json.dump({"channel": ch, "camera": cam, "shot_id": f"{v:05d}"}, f)
            files.append({"id": fid, "file": f"{fid}.png"})
            views.append((ch, cam, vp_inputs["maps"][ch][v]))
    with open(scan_dir / "files.json", "w") as f:
        json.dump({"filesets": [{"id": "PointCloud_1", "files": [{"id": "PointCloud", "file": "PointCloud.ply"}]},
                                {"id": "Segmentation2D_1", "files": files}]}, f)
    task = tasks.SegmentedPointCloud(str(scan_dir), use_colmap_poses=False)
    out = task.run()
    labels = sorted(LABELS)
    want_l, _ = oracle_score(pts, views, labels)
    got_pts, _, col = tasks.read_ply(out.path())
    assert np.array_equal(got_pts, pts)
    assert out.get_metadata("labels") == [labels[i] for i in want_l]
    palette = np.array([tasks.DEFAULT_COLORS[l] for l in labels])
    assert np.array_equal(col, np.round(palette[want_l] * 255).astype(np.uint8))
    assert os.path.basename(out.path()) == "SegmentedPointCloud.ply"
    # colmap poses requested but absent: every view is skipped, like the reference's warning path
    l, s = tasks.SegmentedPointCloud(str(scan_dir), use_colmap_poses=True).compute()[2:5:2]
    assert (s == 0).all() and (l == 0).all()

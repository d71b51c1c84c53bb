"""Volume -> point-cloud handoff: ``plant3dvision_b200.proc3d`` against the oracle
(``oracle/vol2pcd_oracle.py`` = the reference's proc3d.py:490-570 on scipy).

Bars.  Index work is bit-exact: the SET and ORDER of emitted points must equal the reference's
(np.nonzero order of the level-set band, zero-gradient points dropped), and the signed distance must be
bit-identical wherever it is within the transform's cap.  Points and normals are float64 results of
the same operations in the same order; the only place the reference's arithmetic is not pinned is
``np.linalg.norm`` (BLAS ddot, may fuse multiply-adds), so they are gated at 1e-12 absolute on unit
normals / 1e-12 relative on points instead of bit equality (and reported when bit-equal).
"""
import numpy as np
import pytest

from conftest import split_cam, vp_shape_origin


# ---------------------------------------------------------------------------------------------
# CPU: oracle self-consistency and host logic
# ---------------------------------------------------------------------------------------------
def blob(shape, seed=0, noise=0.01):
    rng = np.random.default_rng(seed)
    ii, jj, kk = np.meshgrid(*[np.arange(n) for n in shape], indexing="ij")
    c = [n / 2.0 for n in shape]
    r = min(shape) / 3.0
    vol = (((ii - c[0]) ** 2 + 1.4 * (jj - c[1]) ** 2 + 0.7 * (kk - c[2]) ** 2) < r * r).astype(np.float32)
    vol[rng.random(shape) < noise] = 1.0
    return vol


def test_oracle_loop_and_vectorised_forms_agree():
    from oracle import vol2pcd_oracle as vo
    vol = blob((19, 14, 23))
    p0, n0 = vo.vol2pcd(vol, np.array([1.0, -2.0, 3.0]), 0.5, 1.0, loop=True)
    p1, n1 = vo.vol2pcd(vol, np.array([1.0, -2.0, 3.0]), 0.5, 1.0, loop=False)
    assert p0.shape == p1.shape and len(p0) > 100
    np.testing.assert_allclose(p0, p1, rtol=1e-13, atol=1e-13)
    np.testing.assert_allclose(n0, n1, rtol=0, atol=1e-13)


def test_gaussian_weights_are_scipys():
    from scipy.ndimage import gaussian_filter
    from plant3dvision_b200.proc3d import gaussian_weights
    w, r = gaussian_weights(1.0)
    assert r == 4 and len(w) == 9
    imp = np.zeros(21)
    imp[10] = 1.0
    np.testing.assert_array_equal(gaussian_filter(imp, 1)[6:15], w)


def test_log_cutoff_selects_what_exp_then_threshold_selects():
    from plant3dvision_b200.proc3d import _log_cutoff
    for dt in (np.float32, np.float64):
        c = dt(_log_cutoff(dt))
        assert np.exp(c) > 0.5 and not (np.exp(np.nextafter(c, dt(-np.inf))) > 0.5)
        v = np.linspace(-0.7, -0.68, 20001).astype(dt)
        np.testing.assert_array_equal(v >= c, np.exp(v) > 0.5)


def test_index_point_convention():
    # the reference pins this in tests/unit/test_proc3d.py:12-30
    from plant3dvision_b200.proc3d import index2point, point2index
    idx = np.array([[0, 0, 0], [1, 2, 3]])
    pts = index2point(idx, np.array([0.5, 0.5, 0.5]), 2.0)
    np.testing.assert_array_equal(pts, [[0.5, 0.5, 0.5], [2.5, 4.5, 6.5]])
    np.testing.assert_array_equal(point2index(pts, np.array([0.5, 0.5, 0.5]), 2.0), idx)


def test_multiclass_oracle_matches_hand_case():
    from oracle import vol2pcd_oracle as vo
    vox = {"background": np.full((2, 2, 2), 0.01), "leaf": np.full((2, 2, 2), 0.5), "stem": np.full((2, 2, 2), 0.02)}
    vox["stem"][0, 0, 0] = 0.9
    vox["leaf"][0, 0, 0] = 0.05  # contrast 1/0.05 = 20 > 10 -> stem kept there
    out = vo.multiclass_select(vox)
    assert out["stem"][0, 0, 0] and out["stem"].sum() == 1
    # leaf is the argmax everywhere else and 1 > 10 * max(0.01, 0.02) holds there
    assert out["leaf"].sum() == 7 and not out["leaf"][0, 0, 0]
    # with a contrast of 60 the stem voxel (1 > 60 * 0.05 fails) and the leaf voxels (1 > 60 * 0.02 fails) go
    out = vo.multiclass_select(vox, min_contrast=60.0)
    assert not out["stem"].any() and not out["leaf"].any()


# ---------------------------------------------------------------------------------------------
# GPU parity
# ---------------------------------------------------------------------------------------------
def check_against_oracle(vol, origin, vs, lsv, got=None, oracle_vol=None):
    from oracle import vol2pcd_oracle as vo
    from plant3dvision_b200 import proc3d
    if got is None:
        got = proc3d.vol2pcd(vol, origin, vs, lsv, return_distance=True)
    ovol = vol if oracle_vol is None else oracle_vol
    want_p, want_n = vo.vol2pcd(ovol, origin, vs, lsv)
    # signed distance: bit-identical inside the cap
    sd = vo.signed_distance(ovol)
    near = np.abs(sd) <= got.distance_cap - 0.5
    np.testing.assert_array_equal(got.signed_distance[near], sd[near])
    assert (np.abs(got.signed_distance[~near]) >= got.distance_cap - 0.5).all()
    assert np.array_equal(np.sign(got.signed_distance - 0.25), np.sign(sd - 0.25))
    # same points, same order
    assert got.points.shape == want_p.shape, (got.points.shape, want_p.shape)
    if len(want_p):
        np.testing.assert_allclose(got.points, want_p, rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(got.normals, want_n, rtol=0, atol=1e-12)
    return got, bool(np.array_equal(got.points, want_p) and np.array_equal(got.normals, want_n))


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(19, 14, 23), (40, 33, 70), (8, 8, 8), (2, 2, 2), (3, 50, 2), (64, 64, 64)])
@pytest.mark.parametrize("lsv", [0, 1.0, -0.5])
def test_vol2pcd_blob(shape, lsv):
    vol = blob(shape, seed=sum(shape))
    got, exact = check_against_oracle(vol, np.array([10.0, -3.5, 0.25]), 0.3, lsv)
    if min(shape) > 4:
        assert len(got) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.bool_, np.uint8, np.int32, np.float32, np.float64])
def test_vol2pcd_dtypes_and_border_contact(dtype):
    rng = np.random.default_rng(3)
    vol = rng.random((30, 28, 45)) < 0.2          # noise: object touches every face of the volume
    vol[5:20, :, 10:30] = True                    # slab crossing two faces
    v = vol.astype(dtype)
    if dtype == np.float32:
        v = (v * 0.7 + 0.1).astype(np.float32)    # 0.8 / 0.1 around the 0.5 threshold
    if dtype == np.int32:
        v = np.where(vol, 1, -1).astype(np.int32)  # carving values
    check_against_oracle(v, np.zeros(3), 1.0, 1.0)


@pytest.mark.gpu
def test_vol2pcd_degenerate_volumes_give_no_points():
    from plant3dvision_b200 import proc3d
    for fill in (0.0, 1.0):
        out = proc3d.vol2pcd(np.full((12, 9, 17), fill, dtype=np.float32), np.zeros(3), 1.0, 1.0)
        assert out.points.shape == (0, 3) and out.normals.shape == (0, 3)
    with pytest.raises(ValueError):
        proc3d.vol2pcd(np.zeros((1, 9, 17), dtype=np.float32), np.zeros(3), 1.0, 1.0)  # np.gradient raises too
    with pytest.raises(ValueError):
        proc3d.vol2pcd(np.zeros((9, 17), dtype=np.float32), np.zeros(3), 1.0, 1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("vs", [0.6, 0.3])
def test_handoff_from_live_carving_handle(vp_inputs, vs):
    """Voxels -> PointCloud without leaving HBM == the reference's get_values -> vol2pcd."""
    from plant3dvision_b200.cl import Backprojection
    from plant3dvision_b200 import proc3d
    shape, origin = vp_shape_origin(vp_inputs["bbox"], vs)
    bp = Backprojection(shape, origin, vs, type="carving")
    maps = vp_inputs["maps"]["background"]
    bp.process_views([(*split_cam(vp_inputs["cams"][v]), maps[v]) for v in range(len(maps))], invert=True)
    vol = bp.get_values().copy()
    got = proc3d.vol2pcd(bp, np.asarray(origin, dtype=np.float64), vs, 1.0, return_distance=True)
    bp.close()
    assert (vol == 1).any()
    g, exact = check_against_oracle(vol, np.asarray(origin, dtype=np.float64), vs, 1.0, got=got)
    assert len(g) > 1000


@pytest.mark.gpu
@pytest.mark.parametrize("log", [False, True])
def test_handoff_from_live_averaging_handle(vp_inputs, log):
    from plant3dvision_b200.cl import Backprojection
    from plant3dvision_b200 import proc3d
    vs = 0.6
    shape, origin = vp_shape_origin(vp_inputs["bbox"], vs)
    bp = Backprojection(shape, origin, vs, type="averaging", log=log)
    # log: product of per-view probabilities, so use the (inverted) background channel = plant silhouette,
    # which reaches 1.0; plain averaging: a few views of a thin organ's probability map
    maps = vp_inputs["maps"]["background" if log else "stem"]
    views = [(*split_cam(vp_inputs["cams"][v]), maps[v]) for v in range(0, 18, 3 if log else 5)]
    bp.process_views(views, invert=log)
    vol = bp.get_values().copy()
    assert 0 < (np.exp(vol) > 0.5 if log else vol > 0.5).sum() < vol.size
    got = proc3d.vol2pcd(bp, np.asarray(origin, dtype=np.float64), vs, 0, return_distance=True)
    bp.close()
    ref_vol = vol
    if log:  # tasks/cl.py:172-174
        ref_vol = np.exp(vol)
        ref_vol[ref_vol > 1] = 1.0
    check_against_oracle(None, np.asarray(origin, dtype=np.float64), vs, 0, got=got, oracle_vol=ref_vol)


@pytest.mark.gpu
def test_multiclass_pointcloud_matches_reference_task_logic():
    from oracle import vol2pcd_oracle as vo
    from plant3dvision_b200 import proc3d
    rng = np.random.default_rng(11)
    shape = (24, 20, 36)
    labels = ["background", "leaf", "stem", "fruit"]
    vox = {}
    for q, l in enumerate(labels):
        base = rng.random(shape) * 0.05
        if l != "background":
            base += 0.9 * blob(shape, seed=q, noise=0.0) * (np.roll(blob(shape, seed=q, noise=0.0), 5 * q, axis=2))
        vox[l] = base
    vox["background"] += 0.02
    origin = np.array([0.0, 1.0, 2.0])
    for kw in ({}, {"min_contrast": 1.0, "min_score": 0.0, "background_prior": 0.5}):
        sel = vo.multiclass_select(vox, **{"background_prior": 1.0, "min_contrast": 10.0, "min_score": 0.2, **kw})
        want_p, want_n, want_l = [np.zeros((0, 3))], [np.zeros((0, 3))], []
        for l in labels[1:]:
            p, n = vo.vol2pcd(sel[l], origin, 0.5, 1.0)
            want_p.append(p)
            want_n.append(n)
            want_l += [l] * len(p)
        got, got_l = proc3d.pointcloud_from_voxels(vox, origin, 0.5, 1.0, **kw)
        assert got_l == want_l
        np.testing.assert_allclose(got.points, np.concatenate(want_p), rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(got.normals, np.concatenate(want_n), rtol=0, atol=1e-12)
        assert len(got_l) > 0


@pytest.mark.gpu
def test_vol2pcd_c2_size_properties():
    """Full C2 grid (301 x 301 x 551): properties that need no CPU pass over the volume — every point
    lies within sqrt(3) voxels of a band voxel centre, normals are unit, and the count is stable
    under re-running."""
    from plant3dvision_b200 import proc3d, synth
    from plant3dvision_b200.cl import Backprojection
    w = synth.workload("C2", scale=1.0)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"], device="cuda").cpu().numpy()
    cams = w["cams"].astype(np.float32)
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving")
    bp.process_views([(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], masks[v]) for v in range(len(cams))])
    a = proc3d.vol2pcd(bp, None, w["voxel_size"], 1.0)
    b = proc3d.vol2pcd(bp, None, w["voxel_size"], 1.0)
    bp.close()
    assert len(a) > 10000 and np.array_equal(a.points, b.points) and np.array_equal(a.normals, b.normals)
    np.testing.assert_allclose(np.linalg.norm(a.normals, axis=1), 1.0, rtol=0, atol=1e-12)
    idx = (a.points - np.asarray(w["origin"])) / w["voxel_size"]
    assert (idx > -3).all() and (idx < np.asarray(w["shape"]) + 3).all()


@pytest.mark.gpu
def test_pointcloud_task_on_a_scan_directory(tmp_path):
    """Voxels output on disk (TIFF volume / multi-label NPZ + {origin, voxel_size}) -> PointCloud.ply, as
    tasks/proc3d.py:66-136 does, against the oracle run on the same files."""
    from oracle import vol2pcd_oracle as vo
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200.tasks import proc3d as tp
    from plant3dvision_b200.tasks.cl import write_volume
    from test_host_logic import make_scan
    scan = fsdb_lite.Scan(make_scan(tmp_path))
    origin, vs = [1.0, 2.0, 3.0], 0.5
    # single volume, carving values, TIFF
    vol = np.where(blob((24, 20, 31), seed=1) > 0.5, 1, -1).astype(np.int32)
    fs = scan.get_fileset("Voxels_0123456789", create=True)
    f = fs.create_file("Voxels", "tiff")
    write_volume(f.path(), vol)
    f.set_metadata({"voxel_size": vs, "origin": origin})
    out = tp.PointCloud(scan).run()
    pts, nrm, col = tp.read_ply(out.path())
    want_p, want_n = vo.vol2pcd(vol, np.array(origin), vs, 1.0)
    assert col is None and pts.shape == want_p.shape and len(pts) > 100
    np.testing.assert_allclose(pts, want_p, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(nrm, want_n, rtol=0, atol=1e-12)
    assert out.get_metadata("voxel_size") == vs
    # multiclass NPZ
    rng = np.random.default_rng(2)
    vox = {"background": rng.random((24, 20, 31)) * 0.02 + 0.01}
    for q, l in enumerate(("leaf", "stem")):
        vox[l] = rng.random((24, 20, 31)) * 0.02 + 0.9 * np.roll(blob((24, 20, 31), seed=3, noise=0.0), 6 * q, axis=2)
    f2 = scan.get_fileset("Voxels_multi", create=True).create_file("Voxels", "npz")
    np.savez_compressed(f2.path(), **vox)
    f2.set_metadata({"voxel_size": vs, "origin": origin})
    out2 = tp.PointCloud(scan, voxels_file=f2, level_set_value=0.0).run()
    pts, nrm, col = tp.read_ply(out2.path())
    sel = vo.multiclass_select(vox)
    parts = [vo.vol2pcd(sel[l], np.array(origin), vs, 0.0) for l in ("leaf", "stem")]
    want_p = np.concatenate([p for p, _ in parts])
    assert pts.shape == want_p.shape and len(pts) > 100 and col is not None
    np.testing.assert_allclose(pts, want_p, rtol=1e-12, atol=1e-12)
    labels = out2.get_metadata("labels")
    assert labels == ["leaf"] * len(parts[0][0]) + ["stem"] * len(parts[1][0])
    assert (col[:len(parts[0][0])] == [0, 255, 255]).all() and (col[len(parts[0][0]):] == [0, 0, 255]).all()

"""The mask producer: ``plant3dvision_b200.proc2d.masks`` (bp_masks_filter, one batched GPU call) against
``oracle/masks_oracle.py`` (= tasks/proc2d.py:224-237 with proc2d.py:69-220 on numpy / scipy).

Bar: the output is a binary image, and every float64 operation before the threshold is done separately and in the
reference's order, so masks must be EQUAL byte for byte.
"""
import json

import numpy as np
import pytest

CROSS = np.array([[0, 1, 0], [1, 1, 1], [0, 1, 0]], dtype=np.uint8)
SQUARE = np.ones((3, 3), dtype=np.uint8)
T_UP = np.array([[0, 1, 0], [1, 1, 1], [0, 0, 0]], dtype=np.uint8)       # asymmetric (skimage's "t-shaped" elements)
CORNER = np.array([[1, 0, 0], [0, 0, 0], [0, 0, 1]], dtype=np.uint8)     # no centre, two diagonal taps
RIGHT = np.array([[0, 0, 0], [0, 0, 1], [0, 0, 0]], dtype=np.uint8)      # a pure shift


def plantish(rng, H, W, C=3):
    """Green blobs on a dim noisy background, uint8."""
    yy, xx = np.mgrid[0:H, 0:W]
    img = rng.integers(0, 90, (H, W, C)).astype(np.float64)
    for _ in range(6):
        cy, cx, r = rng.uniform(0, H), rng.uniform(0, W), rng.uniform(2, max(3, min(H, W) / 5))
        blob = ((yy - cy) ** 2 + (xx - cx) ** 2) < r * r
        img[blob, 1] += rng.uniform(60, 160)
        img[blob, 0] += rng.uniform(0, 60)
    return np.clip(img, 0, 255).astype(np.uint8)


# ---------------------------------------------------------------------------------------------
# CPU: the oracle itself, and the dilation semantics the C ABI documents
# ---------------------------------------------------------------------------------------------
def test_oracle_rescale_and_indices_by_hand():
    from oracle import masks_oracle as mo
    img = np.array([[[10, 110, 10], [60, 10, 60]]], dtype=np.uint8)  # min 10, max 110
    lin = mo.linear(img, [0, 1, 0])
    assert lin.dtype == np.float64 and np.array_equal(lin, [[1.0, 0.0]])
    lin2 = mo.linear(img, [0.5, 0.25, 1])
    assert np.array_equal(lin2, [[0.5 * 0.0 + 0.25 * 1.0 + 1 * 0.0, 0.5 * 0.5 + 0.25 * 0.0 + 1 * 0.5]])
    eg = mo.excess_green(img)
    s0, s1 = (0.0 + 1.0) + 0.0 + 1e-9, (0.5 + 0.0) + 0.5 + 1e-9
    assert np.array_equal(eg, [[2 * (1.0 / s0) - 0.0 / s0 - 0.0 / s0, 2 * (0.0 / s1) - 0.5 / s1 - 0.5 / s1]])
    flat = np.full((2, 2, 3), 37, dtype=np.uint8)  # imin == imax: clipped to the output range
    assert np.array_equal(mo.linear(flat, [0, 1, 0]), np.ones((2, 2)))
    assert np.array_equal(mo.masks_filter(img, "linear", [0, 1, 0], 0.3), [[255, 0]])
    with pytest.raises(Exception, match="Unknown masking type"):
        mo.masks_filter(img, "ndvi")


def shift_dilate(img, fp):
    """The formula include/masks_b200.h documents: out[y][x] = OR over set (a, b) of in[y-(a-1)][x-(b-1)]."""
    H, W = img.shape
    pad = np.zeros((H + 2, W + 2), dtype=bool)
    pad[1:-1, 1:-1] = img
    out = np.zeros((H, W), dtype=bool)
    for a in range(3):
        for b in range(3):
            if fp[a][b]:
                out |= pad[1 - (a - 1):1 - (a - 1) + H, 1 - (b - 1):1 - (b - 1) + W]
    return out


@pytest.mark.parametrize("fp", [CROSS, SQUARE, T_UP, CORNER, RIGHT])
def test_documented_dilation_formula_is_scipys(fp):
    from scipy import ndimage as ndi
    from oracle import masks_oracle as mo
    rng = np.random.default_rng(1)
    img = rng.random((23, 41)) < 0.05
    img[0, 0] = img[-1, -1] = img[0, -1] = True
    assert np.array_equal(shift_dilate(img, fp), ndi.binary_dilation(img, structure=fp, border_value=0))
    twice = shift_dilate(shift_dilate(img, fp), fp)
    assert np.array_equal(twice, mo.binary_dilation_sequence(img, ((fp, 2),)))


def test_masks_params_parse_the_reference_configs():
    from plant3dvision_b200.tasks import proc2d as tp
    p = tp.parse_params({"type": "linear", "parameters": "[0,1,0]", "dilation": 1, "threshold": 0.2, "binarize": True},
                        strict=False)  # configs/geom_pipe_full.toml:33-38
    assert p["parameters"] == [0, 1, 0] and p["dilation"] == 1 and p["threshold"] == 0.2 and p["upstream_task"] == "Undistorted"
    assert tp.parse_params({})["type"] == "linear" and tp.parse_params({})["threshold"] == 0.3
    with pytest.raises(TypeError):
        tp.parse_params({"bogus": 1})
    with pytest.raises(Exception, match="Unknown masking type"):
        tp.Masks("/nonexistent", type="ndvi")


# ---------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------
SEQS = {
    "none": (),
    "cross1": ((CROSS, 1),),
    "disk_like": ((T_UP, 1), (T_UP[::-1].copy(), 1), (T_UP.T.copy(), 1), (T_UP.T[:, ::-1].copy(), 1), (CROSS, 2), (SQUARE, 1)),
    "asym": ((CORNER, 2), (RIGHT, 3), (T_UP, 1)),
    "zero_repeat": ((SQUARE, 0), (CROSS, 1)),
}


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(37, 70, 3), (64, 64, 3), (5, 1, 3), (1, 33, 4), (270, 360, 4), (1080, 1440, 3)])
@pytest.mark.parametrize("kind", ["linear", "excess_green"])
def test_masks_equal_oracle(shape, kind):
    from oracle import masks_oracle as mo
    from plant3dvision_b200 import proc2d
    rng = np.random.default_rng(shape[0] * 7 + shape[1])
    H, W, C = shape
    imgs = [plantish(rng, H, W, C) for _ in range(3)]
    imgs.append(np.full((H, W, C), 200, dtype=np.uint8))                      # flat image: imin == imax
    imgs.append(rng.integers(0, 256, (H, W, C), dtype=np.uint8))              # noise: many values near the threshold
    names = list(SEQS) if H * W < 100000 else ["disk_like"]
    for name in names:
        for params, thr in (([0, 1, 0], 0.3), ([0.3, 0.9, -0.4], 0.15)) if kind == "linear" else (([0, 1, 0], 0.0), ([0, 1, 0], 0.2)):
            got = proc2d.masks(imgs, kind, params, thr, SEQS[name])
            for im, g in zip(imgs, got):
                want = mo.masks_filter(im, kind, params, thr, SEQS[name])
                assert g.dtype == np.uint8 and g.shape == (H, W)
                assert np.array_equal(g, want), (name, kind, thr, int((g != want).sum()))


@pytest.mark.gpu
def test_masks_arguments_and_empty_batches():
    from plant3dvision_b200 import proc2d
    assert proc2d.masks([]) == []
    rng = np.random.default_rng(0)
    a = plantish(rng, 20, 30)
    with pytest.raises(Exception, match="Unknown masking type"):
        proc2d.masks([a], "ndvi")
    with pytest.raises(TypeError):
        proc2d.masks([a.astype(np.float32)])
    with pytest.raises(TypeError):
        proc2d.masks([a[:, :, 0]])
    with pytest.raises(ValueError):
        proc2d.masks([a, a[:10]])
    with pytest.raises(ValueError):
        proc2d.masks([a], dilation=((np.ones((5, 5)), 1),))
    assert np.array_equal(proc2d.masks([a], dilation=1)[0], proc2d.masks([a], dilation=((CROSS, 1),))[0])


@pytest.mark.gpu
def test_masks_stay_on_the_device_for_carving(vp_inputs):
    """Masks -> Voxels without the PNG round trip: the device buffer of bp_masks_filter feeds the carve directly."""
    from conftest import vp_shape_origin
    from plant3dvision_b200 import proc2d
    from plant3dvision_b200.cl import Backprojection
    # synthetic RGB renderings whose green channel is the inverted background probability of the fixture
    bg = vp_inputs["maps"]["background"]
    n = 6
    imgs = []
    for v in range(n):
        im = np.zeros((*bg[v].shape, 3), dtype=np.uint8)
        im[:, :, 1] = 255 - bg[v]
        im[:, :, 0] = bg[v] // 3
        imgs.append(im)
    host = proc2d.masks(imgs, "linear", [0, 1, 0], 0.3, 1)
    ptr, n_out, H, W, ms = proc2d.masks(imgs, "linear", [0, 1, 0], 0.3, 1, keep_on_device=True)
    assert (n_out, H, W) == (n, *bg[0].shape) and ptr and ms > 0
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.6)
    cams = vp_inputs["cams"][:n].astype(np.float32)
    a = Backprojection(shape, origin, 0.6, type="carving")
    a.process_views([(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], host[v]) for v in range(n)])
    want = a.get_values().copy()
    b = Backprojection(shape, origin, 0.6, type="carving")
    b.process_device_views(cams, [ptr + v * H * W for v in range(n)], [(H, W)] * n)
    got = b.get_values().copy()
    proc2d.device_free(ptr)
    assert np.array_equal(got, want) and (want == 1).any() and (want == -1).any()
    a.close()
    b.close()


@pytest.mark.gpu
def test_masks_task_on_a_scan_directory(tmp_path):
    from PIL import Image
    from oracle import masks_oracle as mo
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200.tasks import proc2d as tp
    rng = np.random.default_rng(5)
    scan_dir = tmp_path / "scan"
    (scan_dir / "images").mkdir(parents=True)
    (scan_dir / "metadata" / "images").mkdir(parents=True)
    files, imgs = [], {}
    for v in range(5):
        fid = f"{v:05d}_rgb"
        imgs[fid] = plantish(rng, 90, 120)
        Image.fromarray(imgs[fid]).save(scan_dir / "images" / f"{fid}.png")
        with open(scan_dir / "metadata" / "images" / f"{fid}.json", "w") as f:
            json.dump({"channel": "rgb" if v != 3 else "depth", "shot_id": f"{v:05d}"}, f)
        files.append({"id": fid, "file": f"{fid}.png"})
    with open(scan_dir / "files.json", "w") as f:
        json.dump({"filesets": [{"id": "images", "files": files}]}, f)
    out_fs = tp.run_masks(str(scan_dir), {"upstream_task": "ImagesFilesetExists", "type": "linear", "parameters": "[0,1,0]",
                                          "threshold": 0.2, "dilation": 1, "query": {"channel": "rgb"}})
    got = fsdb_lite.Scan(str(scan_dir)).get_fileset(out_fs.id)
    assert [f.id for f in got.get_files()] == [f"{v:05d}_rgb" for v in (0, 1, 2, 4)]  # the query drops file 3
    for f in got.get_files():
        want = mo.masks_filter(imgs[f.id], "linear", [0, 1, 0], 0.2, mo.disk_sequence(1))
        assert np.array_equal(f.read_image(), want) and want.any() and not want.all()
        md = f.get_metadata("Masks")
        assert md == {"upstream_task": "ImagesFilesetExists", "filter": "linear", "threshold": 0.2, "dilation": 1,
                      "linear_coeff": [0, 1, 0], "query": {"channel": "rgb"}}

"""The oracles of the SURVEY §8f rows pinned to the REFERENCE'S OWN TEXT: tests/golden/f_rows_golden.npz holds what
`vol2pcd`, `index2point`, `backproject_points` (plant3dvision/proc3d.py:490-570, :28-45, :655-659) and the scoring loop
of `SegmentedPointCloud.run` (plant3dvision/tasks/proc3d.py:203-232) return when their source, cut out of the reference
by tests/golden/make_golden_f.py, is executed as written.  The CPU tests check oracle == fixture (bit for bit: same
numpy / scipy operations in the same order); the GPU tests check the CUDA path against the same fixtures through the
C ABI, at the bars of tests/test_vol2pcd.py / tests/test_score_points.py."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

VOLUMES = ["ball", "ball_ls1", "blobs", "carve_border", "prob"]


@pytest.fixture(scope="module")
def gold():
    z = np.load(os.path.join(GOLDEN, "f_rows_golden.npz"))
    return {k: z[k] for k in z.files if k != "meta"}, json.loads(str(z["meta"]))


def test_fixture_cites_the_reference_lines(gold):
    _, meta = gold
    assert meta["cites"]["vol2pcd"].startswith("plant3dvision/proc3d.py:")
    assert meta["cites"]["scoring"].startswith("plant3dvision/tasks/proc3d.py:")


@pytest.mark.parametrize("name", VOLUMES)
@pytest.mark.parametrize("loop", [False, True])
def test_vol2pcd_oracle_equals_the_reference_text(gold, name, loop):
    from oracle import vol2pcd_oracle as vo
    g, _ = gold
    args = g[f"v2p_{name}_args"]
    pts, nrm = vo.vol2pcd(g[f"v2p_{name}_volume"], args[0:3], float(args[3]), float(args[4]), loop=loop)
    assert pts.shape == g[f"v2p_{name}_points"].shape
    if loop:  # the per-point function as written: bit-identical
        assert np.array_equal(pts, g[f"v2p_{name}_points"]) and np.array_equal(nrm, g[f"v2p_{name}_normals"])
    else:     # vectorised norm (x*x + y*y + z*z) vs np.linalg.norm per point: last-bit differences allowed
        np.testing.assert_allclose(pts, g[f"v2p_{name}_points"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(nrm, g[f"v2p_{name}_normals"], rtol=0, atol=1e-12)


def test_index2point_oracle_equals_the_reference_text(gold):
    from oracle import vol2pcd_oracle as vo
    g, _ = gold
    idx = np.array([[0, 0, 0], [1, 2, 3], [10, 0, 5]], dtype=np.float64)
    assert np.array_equal(vo.index2point(idx, np.array([1.0, -2.0, 0.5]), 0.25), g["index2point"])


def _score_views(gold, vp_inputs):
    g, meta = gold
    labels = meta["score_label_names"]
    views = []
    for v in g["score_view_index"]:
        c = vp_inputs["cams"][v]
        cam = {"camera_model": {"params": c[0:4].tolist() + [0.0] * 4}, "rotmat": c[4:13].reshape(3, 3).tolist(),
               "tvec": c[13:16].tolist()}
        for lab in labels + ["background"]:
            views.append((lab, cam if v != 6 else None, vp_inputs["maps"][lab][v]))
    return g["score_points"], views, labels


def test_score_points_oracle_equals_the_reference_text(gold, vp_inputs):
    from oracle import vol2pcd_oracle as vo
    pts, views, labels = _score_views(gold, vp_inputs)
    lab, scores = vo.score_points(pts, views, labels)
    assert np.array_equal(scores, gold[0]["score_scores"])
    assert np.array_equal(lab, gold[0]["score_labels"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", VOLUMES)
def test_gpu_vol2pcd_equals_the_reference_text(gold, name):
    from plant3dvision_b200 import proc3d
    g, _ = gold
    args = g[f"v2p_{name}_args"]
    pcd = proc3d.vol2pcd(g[f"v2p_{name}_volume"], args[0:3], float(args[3]), float(args[4]))
    assert pcd.points.shape == g[f"v2p_{name}_points"].shape
    np.testing.assert_allclose(pcd.points, g[f"v2p_{name}_points"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(pcd.normals, g[f"v2p_{name}_normals"], rtol=0, atol=1e-12)


@pytest.mark.gpu
def test_gpu_score_points_equals_the_reference_text(gold, vp_inputs):
    from plant3dvision_b200 import proc3d
    pts, views, labels = _score_views(gold, vp_inputs)
    lab, scores, _ = proc3d.score_points(pts, views, labels)
    assert np.array_equal(scores, gold[0]["score_scores"])
    assert np.array_equal(lab, gold[0]["score_labels"])

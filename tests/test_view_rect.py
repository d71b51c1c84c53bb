"""bp_view_rect (host only): the rectangle of a view that the grid projects into must be CONSERVATIVE under the
reference's own float32 projection (kernels/backprojection.c:3-34) - with BP_VIEW_RECT=1 the library neither copies
nor packs the pixels outside it, so a voxel that sampled one would read garbage.

The check uses the oracle as the reference does its work: carve with a mask that is 1 inside the rectangle and 0
outside.  A voxel that samples a pixel outside the rectangle is carved (-1); none may be.
"""
import numpy as np
import pytest

from oracle.backprojection_oracle import BackprojectionOracle
from plant3dvision_b200 import synth
from plant3dvision_b200.cl import Backprojection


def _carved_outside(shape, origin, vs, cam, H, W, planes=None):
    x0, y0, x1, y1 = Backprojection.view_rect(shape, origin, vs, cam, H, W, planes=planes)
    assert 0 <= x0 < x1 <= W and 0 <= y0 < y1 <= H
    assert x0 % 32 == 0 and y0 % 32 == 0 and (x1 % 32 == 0 or x1 == W) and (y1 % 32 == 0 or y1 == H)
    mask = np.zeros((H, W), dtype=np.uint8)
    mask[y0:y1, x0:x1] = 255
    ora = BackprojectionOracle(shape, origin, vs, type="carving")
    ora.process_view(cam[0:4], cam[4:13], cam[13:16], mask)
    vol = ora.get_values()
    if planes is not None:
        vol = vol[planes[0]:planes[1]]
    return int((vol == -1).sum()), int((vol == 1).sum()), (x0, y0, x1, y1)


@pytest.mark.parametrize("name", ["C2", "C4", "C5"])
def test_rect_is_conservative_on_the_bench_workloads(name):
    w = synth.workload(name, scale=0.06 if name == "C5" else 0.15)
    cams = np.asarray(w["cams"], dtype=np.float32)
    area = []
    for v in range(0, len(cams), max(1, len(cams) // 12)):
        carved, seen, r = _carved_outside(w["shape"], w["origin"], w["voxel_size"], cams[v], w["H"], w["W"])
        assert carved == 0, f"{name} view {v}: {carved} voxels sample pixels outside {r}"
        assert seen > 0
        area.append((r[2] - r[0]) * (r[3] - r[1]) / (w["H"] * w["W"]))
    if name == "C5":  # the point of the exercise: the 1024^3 grid fills ~60 % of a 6000 x 4000 view
        assert np.mean(area) < 0.8


def test_rect_is_conservative_for_random_cameras_and_slabs():
    rng = np.random.default_rng(7)
    H, W = 210, 330
    n_trimmed = 0
    for trial in range(60):
        shape = [int(rng.integers(3, 40)), int(rng.integers(3, 40)), int(rng.integers(3, 70))]
        vs = float(rng.uniform(0.2, 2.0))
        origin = rng.uniform(-30, 30, 3).astype(np.float32)
        centre = origin + 0.5 * vs * np.asarray(shape)
        # a camera somewhere around the grid (sometimes inside it or very near: the whole image must come back)
        dist = float(rng.choice([0.3, 1.0, 2.0, 5.0, 12.0])) * vs * max(shape)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        pos = centre + dist * d
        look = centre + rng.normal(size=3) * vs * max(shape) * 0.4 - pos
        zc = look / np.linalg.norm(look)
        up = rng.normal(size=3)
        xc = np.cross(up, zc)
        xc /= np.linalg.norm(xc)
        yc = np.cross(zc, xc)
        R = np.stack([xc, yc, zc])
        t = -R @ pos
        f = float(rng.uniform(80, 900))
        cam = np.concatenate([[f, f * rng.uniform(0.9, 1.1), W / 2 + rng.uniform(-40, 40), H / 2 + rng.uniform(-40, 40)],
                              R.reshape(-1), t]).astype(np.float32)
        planes = None
        if trial % 3 == 0:
            b = int(rng.integers(0, shape[0]))
            planes = (b, int(rng.integers(b + 1, shape[0] + 1)))
        carved, _, r = _carved_outside(shape, origin, vs, cam, H, W, planes=planes)
        assert carved == 0, f"trial {trial}: {carved} voxels sample pixels outside {r} (planes {planes})"
        n_trimmed += r != (0, 0, W, H)
    assert n_trimmed > 10  # the test does exercise trimmed rectangles, not only the fallback


def test_rect_fallbacks():
    shape, origin, vs, H, W = [16, 16, 16], [0.0, 0.0, 0.0], 1.0, 100, 200
    eye = np.eye(3, dtype=np.float32).reshape(-1)
    # camera inside the box: a corner is behind the image plane -> the whole image
    cam = np.concatenate([[100, 100, 100, 50], eye, [-8, -8, -8]]).astype(np.float32)
    assert Backprojection.view_rect(shape, origin, vs, cam, H, W) == (0, 0, W, H)
    # non-finite parameters -> the whole image
    cam = np.concatenate([[np.nan, 100, 100, 50], eye, [0, 0, 50]]).astype(np.float32)
    assert Backprojection.view_rect(shape, origin, vs, cam, H, W) == (0, 0, W, H)
    # box entirely outside the picture -> the first tile (never empty)
    cam = np.concatenate([[100, 100, 100, 50], eye, [5000, 0, 50]]).astype(np.float32)
    assert Backprojection.view_rect(shape, origin, vs, cam, H, W) == (0, 0, 32, 32)
    # an empty plane range is an argument error only when reversed; [b, b) gives the whole image
    assert Backprojection.view_rect(shape, origin, vs, cam, H, W, planes=(4, 4)) == (0, 0, W, H)

"""CPU tests of the ORACLE (test infrastructure): the C restatement ``oracle/bp_oracle.c`` against
(1) the committed outputs of the reference's own kernel source (``tests/golden``; produced by
``tests/golden/make_golden.py`` from /root/reference's kernels/backprojection.c compiled unmodified),
(2) that compiled reference source itself (``oracle/_ref``) when it is present, on seeded random
scenes, and (3) an independent numpy restatement of kernels/backprojection.c:3-84.

No GPU, no product code: this is what pins the checker the ``-m gpu`` parity tests rely on.
"""
import hashlib

import numpy as np
import pytest

from conftest import CHANNELS, split_cam, vp_shape_origin
from oracle import backprojection_oracle as bo
from oracle.backprojection_oracle import BackprojectionOracle


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_vp(vp_inputs, kind, vs, channel, invert, log=False, backend="restatement"):
    shape, origin = vp_shape_origin(vp_inputs["bbox"], vs)
    o = BackprojectionOracle(shape, origin, vs, type=kind, log=log, backend=backend)
    for v in range(18):
        m = vp_inputs["maps"][channel][v]
        o.process_view(*split_cam(vp_inputs["cams"][v]), np.invert(m) if invert else m)
    return o.get_values()


# ---------------------------------------------------------------------------------------------
# (1) golden vectors of the reference's kernel source on the reference's virtual_plant fixture
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("vs", [0.6, 0.3])
def test_restatement_carving_matches_reference_golden(vp_inputs, vp_golden, vs):
    vol = run_vp(vp_inputs, "carving", vs, "background", invert=True)
    meta = vp_golden["meta"][f"carve_vs{vs}"]
    assert list(vol.shape) == meta["shape"]
    assert vol.dtype == np.int32
    assert {str(k): int((vol == k).sum()) for k in (-1, 0, 1)} == meta["hist"]
    assert sha(vol) == meta["sha256"]
    assert np.array_equal(vol.astype(np.int8), vp_golden["arrays"][f"carve_vs{vs}"])


@pytest.mark.parametrize("log", [False, True])
@pytest.mark.parametrize("channel", CHANNELS)
def test_restatement_averaging_matches_reference_golden(vp_inputs, vp_golden, channel, log):
    vol = run_vp(vp_inputs, "averaging", 0.6, channel, invert=False, log=log)
    key = f"avg_vs0.6_log{int(log)}_{channel}"
    meta = vp_golden["meta"][key]
    assert vol.dtype == np.float32 and list(vol.shape) == meta["shape"]
    assert sha(vol) == meta["sha256"]  # bit-exact: same float32 operations in the same order
    if key in vp_golden["arrays"]:
        assert np.array_equal(vol, vp_golden["arrays"][key])


def test_lut_path_equals_float_path(vp_inputs):
    """uint8 map + 256-entry table (what the CUDA path uses) == img_as_float32 [+ log] on the host."""
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.6)
    for log in (False, True):
        a = BackprojectionOracle(shape, origin, 0.6, type="averaging", log=log)
        b = BackprojectionOracle(shape, origin, 0.6, type="averaging", log=log)
        for v in range(6):
            m = vp_inputs["maps"]["stem"][v]
            a.process_view(*split_cam(vp_inputs["cams"][v]), m)
            b.process_view_u8_lut(*split_cam(vp_inputs["cams"][v]), m)
        assert np.array_equal(a.get_values(), b.get_values())


# ---------------------------------------------------------------------------------------------
# (2) restatement vs the reference's own kernel source (when oracle/_ref was built here)
# ---------------------------------------------------------------------------------------------
def random_scene(seed, n_views=5, H=97, W=131):
    rng = np.random.default_rng(seed)
    shape = [int(rng.integers(3, 24)), int(rng.integers(3, 24)), int(rng.integers(3, 40))]
    vs = float(rng.uniform(0.5, 3.0))
    origin = [float(x) for x in rng.uniform(-40, 10, 3)]
    views = []
    for _ in range(n_views):
        # a camera somewhere around the grid, sometimes inside it, looking roughly at its centre
        A = rng.normal(size=(3, 3))
        Q, _ = np.linalg.qr(A)
        t = rng.normal(size=3) * 30 + np.array([0, 0, 60.0])
        intr = np.array([rng.uniform(50, 200), rng.uniform(50, 200), W / 2, H / 2], dtype=np.float32)
        mask = (rng.random((H, W)) < 0.6).astype(np.uint8) * rng.integers(1, 256, (H, W)).astype(np.uint8)
        views.append((intr, Q.astype(np.float32).reshape(-1), t.astype(np.float32), mask))
    return shape, origin, vs, views


@pytest.mark.skipif(not bo.have_backend("ref_off"), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("kind", ["carving", "averaging"])
def test_restatement_equals_reference_source(kind, seed):
    shape, origin, vs, views = random_scene(seed)
    a = BackprojectionOracle(shape, origin, vs, type=kind, log=(seed % 2 == 1))
    b = BackprojectionOracle(shape, origin, vs, type=kind, log=(seed % 2 == 1), backend="ref_off")
    for v in views:
        a.process_view(*v)
        b.process_view(*v)
    assert np.array_equal(a.get_values(), b.get_values())
    if kind == "carving":
        assert set(np.unique(a.get_values())) <= {-1, 0, 1}


def test_restatement_slab_offset_equals_full_grid():
    shape, origin, vs, views = random_scene(11)
    full = BackprojectionOracle(shape, origin, vs, type="averaging")
    for v in views:
        full.process_view(*v)
    i0 = shape[0] // 2
    part = BackprojectionOracle([shape[0] - i0, shape[1], shape[2]], origin, vs, type="averaging", i_offset=i0)
    for v in views:
        part.process_view(*v)
    assert np.array_equal(full.get_values()[i0:], part.get_values())


# ---------------------------------------------------------------------------------------------
# (3) independent numpy restatement of kernels/backprojection.c:3-84 (float32 op by op)
# ---------------------------------------------------------------------------------------------
def numpy_backproject(shape, origin, vs, intr, rot, tvec, H, W):
    f = np.float32
    i, j, k = np.meshgrid(*[np.arange(n) for n in shape], indexing="ij")  # common.h:1-12, k fastest
    x = f(origin[0]) + i.astype(f) * f(vs)  # backprojection.c:46-48
    y = f(origin[1]) + j.astype(f) * f(vs)
    z = f(origin[2]) + k.astype(f) * f(vs)
    r, t = rot.astype(f), tvec.astype(f)
    with np.errstate(all="ignore"):
        pz = r[6] * x + r[7] * y + r[8] * z + t[2]  # :11
        px = r[0] * x + r[1] * y + r[2] * z + t[0]  # :17
        py = r[3] * x + r[4] * y + r[5] * z + t[1]  # :18
        u = px / pz * intr[0] + intr[2]  # :20
        v = py / pz * intr[1] + intr[3]  # :21
        ok = ~(pz < 0) & np.isfinite(u) & np.isfinite(v) & (np.abs(u) < 2.0 ** 31) & (np.abs(v) < 2.0 ** 31)
        xi = np.where(ok, np.trunc(np.where(ok, u, 0)), -1).astype(np.int64)  # :23-24 (int) truncation
        yi = np.where(ok, np.trunc(np.where(ok, v, 0)), -1).astype(np.int64)
    ok &= (xi >= 0) & (xi <= W - 1) & (yi >= 0) & (yi <= H - 1)  # :26-31
    return ok, xi, yi


@pytest.mark.parametrize("seed", [3, 4])
def test_restatement_equals_numpy_restatement(seed):
    shape, origin, vs, views = random_scene(seed)
    carve = BackprojectionOracle(shape, origin, vs, type="carving")
    avg = BackprojectionOracle(shape, origin, vs, type="averaging")
    want_c = np.zeros(shape, dtype=np.int32)
    want_a = np.zeros(shape, dtype=np.float32)
    for (intr, rot, tvec, mask) in views:
        carve.process_view(intr, rot, tvec, mask)
        avg.process_view(intr, rot, tvec, mask)
        H, W = mask.shape
        ok, xi, yi = numpy_backproject(shape, origin, vs, intr, rot, tvec, H, W)
        pix = mask[np.clip(yi, 0, H - 1), np.clip(xi, 0, W - 1)]
        live = ok & (want_c != -1)  # backprojection.c:65-67
        want_c = np.where(live & (pix == 0), -1, np.where(live & (pix != 0) & (want_c == 0), 1, want_c))  # :79-83
        val = np.multiply(pix, 1.0 / 255, dtype=np.float32)  # cl.py:205-206
        want_a = np.where(ok, want_a + val, want_a).astype(np.float32)  # :54
    assert np.array_equal(carve.get_values(), want_c)
    assert np.array_equal(avg.get_values(), want_a)


def test_unknown_type_raises():
    with pytest.raises(ValueError):
        BackprojectionOracle([2, 2, 2], [0, 0, 0], 1.0, type="nope")

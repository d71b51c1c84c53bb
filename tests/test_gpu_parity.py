"""GPU parity tests proper: the CUDA path, called through the C ABI (ctypes -> libbp_b200.so),
against (1) the committed outputs of the reference's own kernel source on the reference's
virtual_plant fixture and (2) the CPU oracle on seeded synthetic scans.

Bars: carving bit-exact; averaging bit-exact too on one GPU (the fused kernel sums in view
order), and within 1e-5 relative wherever the summation order changes (multi-GPU tests).
"""
import hashlib

import numpy as np
import pytest

from conftest import CHANNELS, split_cam, vp_shape_origin

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def BP():
    from plant3dvision_b200.cl import Backprojection
    return Backprojection


@pytest.fixture(scope="module")
def Oracle():
    from oracle.backprojection_oracle import BackprojectionOracle
    return BackprojectionOracle


def vp_views(vp_inputs, channel, invert):
    maps = vp_inputs["maps"][channel]
    out = []
    for v in range(maps.shape[0]):
        m = np.invert(maps[v]) if invert else maps[v]
        out.append((*split_cam(vp_inputs["cams"][v]), m))
    return out


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


# ---------------------------------------------------------------------------------------------
# golden vectors: reference kernel source on the reference's virtual_plant fixture
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("vs", [0.6, 0.3, 0.1, 0.06])  # 0.06 = configs/geom_pipe_virtual.toml:6 (BASELINE.json configs[0])
@pytest.mark.parametrize("fused", [True, False])
def test_golden_carving(BP, vp_inputs, vp_golden, vs, fused):
    shape, origin = vp_shape_origin(vp_inputs["bbox"], vs)
    meta = vp_golden["meta"][f"carve_vs{vs}"]
    assert shape == meta["shape"]
    bp = BP(shape, origin, vs, type="carving")
    # geom_pipe_virtual.toml: background channel, invert=True. The device does the inversion.
    views = vp_views(vp_inputs, "background", invert=False)
    if fused:
        bp.process_views(views, invert=True)
    else:
        for (i, r, t, m) in views:
            bp.process_view(i, r, t, np.invert(m))
    vol = bp.get_values()
    assert vol.dtype == np.int32 and list(vol.shape) == shape
    hist = {str(k): int((vol == k).sum()) for k in (-1, 0, 1)}
    assert hist == meta["hist"]
    assert sha(vol) == meta["sha256"]
    key = f"carve_vs{vs}"
    if key in vp_golden["arrays"]:
        assert np.array_equal(vol, vp_golden["arrays"][key].astype(np.int32))
    bp.close()


@pytest.mark.parametrize("vs,log", [(0.6, False), (0.6, True), (0.3, False), (0.3, True)])
def test_golden_averaging_all_labels(BP, vp_inputs, vp_golden, vs, log):
    shape, origin = vp_shape_origin(vp_inputs["bbox"], vs)
    bp = BP(shape, origin, vs, type="averaging", log=log)
    for n, label in enumerate(CHANNELS):
        if n:
            bp.clear()
        bp.process_views(vp_views(vp_inputs, label, invert=False))
        vol = bp.get_values()
        meta = vp_golden["meta"][f"avg_vs{vs}_log{int(log)}_{label}"]
        assert vol.dtype == np.float32 and list(vol.shape) == meta["shape"]
        key = f"avg_vs{vs}_log{int(log)}_{label}"
        if key in vp_golden["arrays"]:
            np.testing.assert_array_equal(vol, vp_golden["arrays"][key])
        assert sha(vol) == meta["sha256"], f"{label}: averaging volume differs from the reference kernel's"
    bp.close()


def test_golden_averaging_streaming_per_view(BP, vp_inputs, vp_golden):
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.3)
    bp = BP(shape, origin, 0.3, type="averaging", log=True)
    for view in vp_views(vp_inputs, "stem", invert=False):
        bp.process_view(*view)
    assert sha(bp.get_values()) == vp_golden["meta"]["avg_vs0.3_log1_stem"]["sha256"]
    bp.close()


# ---------------------------------------------------------------------------------------------
# seeded synthetic scans vs the CPU oracle
# ---------------------------------------------------------------------------------------------
def synth_views(name, scale, soft=0.0, select=None):
    from plant3dvision_b200 import synth
    w = synth.workload(name, scale)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"], soft_px=soft, select=select).numpy()
    return w, [(*split_cam(w["cams"][v]), masks[v]) for v in range(len(w["cams"]))]


def run_both(BP, Oracle, shape, origin, vs, views, kind, log=False, default_value=0, fused=True, invert=False):
    bp = BP(shape, origin, vs, type=kind, log=log, default_value=default_value)
    if fused:
        bp.process_views(views, invert=invert)
    else:
        for (i, r, t, m) in views:
            bp.process_view(i, r, t, np.invert(m) if invert else m)
    got = bp.get_values().copy()
    bp.close()
    ora = Oracle(shape, origin, vs, type=kind, log=log, default_value=default_value)
    for (i, r, t, m) in views:
        ora.process_view(i, r, t, np.invert(m) if invert else m)
    return got, ora.get_values()


@pytest.mark.parametrize("name,scale", [("C2", 0.25), ("C4", 0.125), ("C5", 0.0625)])
def test_synthetic_carving(BP, Oracle, name, scale):
    w, views = synth_views(name, scale)
    got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views, "carving")
    assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ"
    assert (want == 1).sum() > 0 and (want == -1).sum() > 0


@pytest.mark.parametrize("p_zero", [0.01, 0.08, 0.5])
def test_speckle_masks_walk_every_warp_through_every_live_count(BP, Oracle, p_zero):
    """Random salt-and-pepper masks: every view carves a random p_zero of the voxels it sees, so the number of live
    voxels of each warp patch decays 128 -> 0 through every value and in random lane patterns.  This drives the
    dense -> medium (two survivors per lane) -> sparse (g views at a time, re-dealing) hand-offs of carve_kernel
    through all their cases (structured phantoms leave most patches either full or empty)."""
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 0.2)
    rng = np.random.default_rng(int(p_zero * 1000))
    views = [(*split_cam(w["cams"][v]), np.where(rng.random((w["H"], w["W"])) < p_zero, 0, 255).astype(np.uint8))
             for v in range(len(w["cams"]))]
    got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views, "carving")
    assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ"
    live = (want != -1).mean()
    assert (want == -1).any() and (p_zero > 0.05 or live > 0.3)  # slow decay keeps patches above 32 live for long
    # the streaming path (one launch per view: every hand-off restarts from the stored state) agrees too
    got1, want1 = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views[:12], "carving", fused=False)
    assert np.array_equal(got1, want1)


@pytest.mark.parametrize("log", [False, True])
def test_synthetic_averaging(BP, Oracle, log):
    w, views = synth_views("C2", 0.25, soft=2.0)
    got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views, "averaging", log=log)
    np.testing.assert_array_equal(got, want)


def test_ragged_shapes_and_padding(BP, Oracle):
    """nz not a multiple of 4/16/64, ny not a multiple of 16, single-voxel grids."""
    w, views = synth_views("C2", 0.2)
    for shape in ([1, 1, 1], [3, 5, 7], [2, 17, 65], [5, 16, 64], [4, 33, 129], [1, 1, 250]):
        for kind in ("carving", "averaging"):
            got, want = run_both(BP, Oracle, shape, w["origin"], 150.0 / max(shape), views[:7], kind)
            np.testing.assert_array_equal(got, want, err_msg=f"{kind} {shape}")


def test_views_of_different_sizes(BP, Oracle):
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 0.2)
    views = []
    for v, (H, W) in enumerate([(216, 288), (100, 333), (65, 31), (1, 1), (300, 97)]):
        cam = w["cams"][v * 7].copy()
        cam[2], cam[3] = W / 2.0, H / 2.0
        m = synth.render_masks(cam[None], w["capsules"], H, W).numpy()[0]
        m[::3, ::2] = 255
        views.append((*split_cam(cam), m))
    for kind in ("carving", "averaging"):
        got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views, kind)
        np.testing.assert_array_equal(got, want, err_msg=kind)


def test_camera_inside_grid_and_image_borders(BP, Oracle):
    """Voxels behind the camera (p_z < 0), on the camera plane (p_z == 0 -> Inf/NaN), and pixels
    in (-1, 0) which the truncating conversion accepts as column / row 0
    (kernels/backprojection.c:13, :23-31)."""
    rng = np.random.default_rng(3)
    shape, origin, vs = [40, 40, 40], [-20.0, -20.0, -20.0], 1.0
    views = []
    for v in range(6):
        R = np.linalg.qr(rng.normal(size=(3, 3)))[0]
        if np.linalg.det(R) < 0:
            R[0] *= -1
        C = rng.uniform(-10, 10, size=3).round()  # integer camera centre: exact p_z == 0 voxels
        if v == 0:
            R, C = np.eye(3), np.zeros(3)
        t = -R @ C
        H, W = 64, 48
        # principal point near the border so many voxels land in the (-1, 0) band and beyond
        intr = np.array([30.0, 30.0, 0.3 if v % 2 else W / 2, 0.4 if v % 2 else H / 2], dtype=np.float32)
        m = (rng.random((H, W)) < 0.7).astype(np.uint8) * 255
        views.append((intr, R.astype(np.float32).reshape(-1), t.astype(np.float32), m))
    for kind in ("carving", "averaging"):
        got, want = run_both(BP, Oracle, shape, origin, vs, views, kind)
        np.testing.assert_array_equal(got, want, err_msg=kind)
        assert (want == 0).sum() > 0  # some voxels are never seen


@pytest.mark.parametrize("default_value", [0, 5, -1, -3])
def test_carving_default_value(BP, Oracle, default_value):
    """cl.py:173 seeds the grid; kernels/backprojection.c:65-83 only promotes `== 0` and skips `== -1`."""
    w, views = synth_views("C2", 0.15)
    # narrow images so that part of the grid is never seen
    views = [(i, r, t, m[:, : m.shape[1] // 2]) for (i, r, t, m) in views[:5]]
    got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views, "carving",
                         default_value=default_value)
    np.testing.assert_array_equal(got, want)


def test_averaging_default_value(BP, Oracle):
    w, views = synth_views("C2", 0.15, soft=1.5)
    got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], views[:9], "averaging",
                         default_value=0.25)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("kind", ["carving", "averaging"])
@pytest.mark.parametrize("dtype", ["bool", "uint8", "uint16", "int32", "int64", "float32", "float64"])
@pytest.mark.parametrize("invert", [False, True])
def test_mask_dtypes(BP, Oracle, kind, dtype, invert):
    """cl.py:205-215 / :300-301 for every mask dtype io.read_image can plausibly return."""
    w, views = synth_views("C2", 0.15, soft=2.0)
    conv = []
    for (i, r, t, m) in views[:6]:
        if dtype == "bool":
            mm = m > 127
        elif dtype == "uint16":
            mm = m.astype(np.uint16) * 257
        elif dtype.startswith("float"):
            mm = (m / 255.0).astype(dtype) * (3.0 if kind == "carving" else 1.0)
        else:
            mm = m.astype(dtype)
        conv.append((i, r, t, mm))
    if invert and dtype.startswith("float"):
        bp = BP(w["shape"], w["origin"], w["voxel_size"], type=kind)
        with pytest.raises(TypeError):
            bp.process_views(conv, invert=True)
        bp.close()
        return
    if kind == "averaging" and dtype in ("int32", "int64"):
        pytest.skip("signed-int probability maps are not produced by the pipeline")
    got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], conv, kind, log=(kind == "averaging"),
                         invert=invert)
    np.testing.assert_array_equal(got, want)


def test_slabs_equal_full_grid(BP, Oracle):
    """Axis-0 slab handles are bit-identical to the same planes of a full run."""
    w, views = synth_views("C2", 0.25, soft=2.0)
    shape, origin, vs = w["shape"], w["origin"], w["voxel_size"]
    for kind in ("carving", "averaging"):
        full = BP(shape, origin, vs, type=kind, log=(kind == "averaging"))
        full.process_views(views)
        want = full.get_values().copy()
        full.close()
        cuts = [0, 1, shape[0] // 3, shape[0] // 3, shape[0] - 2, shape[0]]
        parts = []
        for a, b in zip(cuts[:-1], cuts[1:]):
            s = BP(shape, origin, vs, type=kind, log=(kind == "averaging"), i_range=(a, b))
            s.process_views(views)
            part = s.get_values().copy()
            assert list(part.shape) == [b - a, shape[1], shape[2]]
            parts.append(part)
            s.close()
        np.testing.assert_array_equal(np.concatenate(parts, axis=0), want)


def test_stripes_equal_full_grid(BP):
    """Striped handles (blocks of 2^lg planes dealt round robin) hold exactly the planes of a full run."""
    w, views = synth_views("C2", 0.25, soft=2.0)
    shape, origin, vs = w["shape"], w["origin"], w["voxel_size"]
    for kind in ("carving", "averaging"):
        full = BP(shape, origin, vs, type=kind, log=(kind == "averaging"))
        full.process_views(views)
        want = full.get_values().copy()
        full.close()
        for lg, world in ((0, 2), (2, 3), (3, 8), (7, 2)):
            got = np.full_like(want, 77)
            seen = np.zeros(shape[0], dtype=int)
            for rank in range(world):
                s = BP(shape, origin, vs, type=kind, log=(kind == "averaging"), stripes=(lg, world, rank))
                assert list(s.local_shape) == [len(s.planes), shape[1], shape[2]]
                s.process_views(views)
                if len(s.planes):
                    got[s.planes] = s.get_values()
                seen[s.planes] += 1
                s.close()
            assert (seen == 1).all()
            np.testing.assert_array_equal(got, want)


def test_view_order_independence_of_carving(BP):
    """The final carve state does not depend on view order - the property the
    view-sharded MIN all-reduce relies on."""
    w, views = synth_views("C2", 0.2)
    a = BP(w["shape"], w["origin"], w["voxel_size"])
    a.process_views(views)
    b = BP(w["shape"], w["origin"], w["voxel_size"])
    b.process_views(views[::-1][:30])
    b.process_views(views[:30][::-1])
    np.testing.assert_array_equal(a.get_values(), b.get_values())
    a.close()
    b.close()


def test_more_views_than_one_launch_holds(BP, Oracle):
    """> 256 views are split over several launches with the state carried in HBM."""
    w, views = synth_views("C2", 0.1)
    many = (views * 5)[:290]
    for kind in ("carving", "averaging"):
        got, want = run_both(BP, Oracle, w["shape"], w["origin"], w["voxel_size"], many, kind)
        np.testing.assert_array_equal(got, want)


# ---------------------------------------------------------------------------------------------
# API behaviour of the drop-in (cl.py:118-311)
# ---------------------------------------------------------------------------------------------
def test_constructs_like_reference_unit_test(BP):
    # tests/unit/test_cl.py:4-9 of the reference
    assert BP([10, 10, 10], [0, 0, 0], 1.0) is not None
    assert BP([10, 10, 10], [0, 0, 0], 1.0, 'averaging') is not None
    with pytest.raises(ValueError):
        BP([10, 10, 10], [0, 0, 0], 1.0, 'median')


def test_get_values_aliases_host_buffer_and_clear(BP):
    w, views = synth_views("C2", 0.1)
    bp = BP(w["shape"], w["origin"], w["voxel_size"], default_value=0)
    v0 = bp.get_values()
    assert v0.dtype == np.int32 and (v0 == 0).all()
    bp.process_views(views[:10])
    v1 = bp.get_values()
    assert np.shares_memory(v0, v1) and np.shares_memory(v1, bp.values_h)  # cl.py:229-232
    assert (v1 != 0).any()
    assert bp.get_volume is not None and np.array_equal(bp.get_volume(), v1)
    bp.clear()
    assert (bp.get_values() == 0).all()
    bp.close()
    with pytest.raises(RuntimeError):
        bp.get_values()


def test_empty_inputs(BP):
    bp = BP([4, 4, 4], [0, 0, 0], 1.0)
    bp.process_views([])
    assert (bp.get_values() == 0).all()
    bp.close()
    z = BP([0, 4, 4], [0, 0, 0], 1.0)
    assert z.get_values().shape == (0, 4, 4)
    z.close()
    bp = BP([4, 4, 4], [0, 0, 0], 1.0)
    with pytest.raises(ValueError):
        bp.process_view(np.ones(4, np.float32), np.eye(3, dtype=np.float32), np.zeros(3, np.float32),
                        np.zeros((4, 4, 3), np.uint8))
    bp.close()


class _F:
    def __init__(self, fid, md, img):
        self.id, self._md, self._img = fid, md, img

    def get_metadata(self, key=None, default=None):
        return self._md if key is None else self._md.get(key, default)

    def read_image(self):
        return self._img


def test_process_fileset_labels_and_skips(BP, Oracle, vp_inputs, caplog):
    """cl.py:234-305: per-label loop with clear(), float64 stack, channel filter, warn+skip on
    missing camera metadata."""
    shape, origin = vp_shape_origin(vp_inputs["bbox"], 0.6)
    files = []
    for v in range(18):
        c = vp_inputs["cams"][v]
        cam = {"camera_model": {"params": list(c[0:4]) + [0, 0, 0, 0]}, "rotmat": c[4:13].reshape(3, 3).tolist(),
               "tvec": c[13:16].tolist()}
        for ch in ("stem", "leaf"):
            md = {"channel": ch, "camera": cam}
            if v == 4 and ch == "stem":
                md = {"channel": ch}  # missing camera -> skipped with a warning
            files.append(_F(f"{v:05d}_{ch}", md, vp_inputs["maps"][ch][v]))
    bp = BP(shape, origin, 0.6, type="averaging", labels=["stem", "leaf"], log=True)
    import logging
    with caplog.at_level(logging.WARNING):
        got = bp.process_fileset(files, "camera")
    assert "skipping" in caplog.text
    ora = Oracle(shape, origin, 0.6, type="averaging", labels=["stem", "leaf"], log=True)
    want = ora.process_fileset(files, "camera")
    assert got.dtype == np.float64 and got.shape == (2, *shape)
    np.testing.assert_array_equal(got, want)
    bp.close()


# ---------------------------------------------------------------------------------------------
# full BASELINE sizes: size-independent properties + oracle on sampled planes
# ---------------------------------------------------------------------------------------------
def test_c2_full_size_properties(BP, Oracle):
    """C2 at its real size (301x301x551 x 60 views @1440x1080): sampled planes equal the oracle run
    on just those planes (slab oracle), carving is idempotent under a second pass of the same views,
    and averaging is linear in the view set (two halves sum to the whole within 1e-5 relative)."""
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 1.0)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"], device="cuda").cpu().numpy()
    views = [(*split_cam(w["cams"][v]), masks[v]) for v in range(60)]
    shape, origin, vs = w["shape"], w["origin"], w["voxel_size"]
    bp = BP(shape, origin, vs)
    bp.process_views(views)
    vol = bp.get_values().copy()
    bp.process_views(views)  # idempotence
    assert np.array_equal(bp.get_values(), vol)
    bp.close()
    for i0 in (0, 150, 299):
        ora = Oracle(shape[:0] + [2] + shape[1:], origin, vs, i_offset=i0)
        for view in views:
            ora.process_view(*view)
        assert np.array_equal(vol[i0:i0 + 2], ora.get_values()), f"planes {i0}.."
    assert 0 < (vol == 1).sum() < 0.05 * vol.size
    # averaging: linearity over view subsets
    tot = BP(shape, origin, vs, type="averaging")
    tot.process_views(views)
    whole = tot.get_values().copy()
    tot.clear()
    tot.process_views(views[:30])
    first = tot.get_values().copy()
    tot.clear()
    tot.process_views(views[30:])
    second = tot.get_values()
    scale = np.abs(whole).max()
    assert np.abs((first + second) - whole).max() <= 1e-5 * scale
    tot.close()


# ---------------------------------------------------------------------------------------------
# multi-GPU plumbing (parallel.py)
# ---------------------------------------------------------------------------------------------
def test_state_tensor_aliases_device_state(BP):
    import torch
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 0.1)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    cams = w["cams"].astype(np.float32)
    views = [(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], masks[v]) for v in range(len(cams))]
    with BP(w["shape"], w["origin"], w["voxel_size"], type="carving") as bp:
        st = bp.state_tensor()
        bp.sync()
        assert st.dtype == torch.int8 and st.shape[0] == w["shape"][0] and int(st.min()) == 2  # all unseen
        bp.process_views(views)
        bp.sync()
        vol = bp.get_values()
        nz = w["shape"][2]
        from plant3dvision_b200.parallel import decode_carve_state
        assert np.array_equal(decode_carve_state(st[:, :, :nz].cpu().numpy()), vol)
    with BP(w["shape"], w["origin"], w["voxel_size"], type="averaging") as bp:
        bp.process_views(views[:3])
        bp.sync()
        assert np.array_equal(bp.state_tensor()[:, :, :w["shape"][2]].cpu().numpy(), bp.get_values())


def test_distributed_world_of_one_equals_plain(BP):
    from plant3dvision_b200 import synth
    from plant3dvision_b200.parallel import DistributedBackprojection
    w = synth.workload("C2", 0.1)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    cams = w["cams"].astype(np.float32)
    views = [(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], masks[v]) for v in range(len(cams))]
    for sharding in ("slabs", "views"):
        with BP(w["shape"], w["origin"], w["voxel_size"]) as a, \
                DistributedBackprojection(w["shape"], w["origin"], w["voxel_size"], sharding=sharding) as b:
            a.process_views(views)
            b.process_views(views)
            assert np.array_equal(a.get_values(), b.get_values())


@pytest.mark.parametrize("kind,sharding", [("carving", "views"), ("carving", "slabs"), ("averaging", "views"),
                                           ("averaging", "slabs"), ("carving", "stripes"), ("averaging", "stripes")])
def test_two_gpus_nccl_equal_sequential(tmp_path, Oracle, kind, sharding):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import socket
    import torch.multiprocessing as mp
    from _gloo_worker import scene, worker
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(worker, args=(2, port, kind, sharding, 0, str(tmp_path), "nccl"), nprocs=2, join=True)
    shape, origin, vs, views = scene()
    o = Oracle(shape, origin, vs, type=kind)
    for v in views:
        o.process_view(*v)
    for rank in range(2):
        vol = np.load(tmp_path / f"rank{rank}.npz")["vol"]
        if kind == "carving" or sharding in ("slabs", "stripes"):
            assert np.array_equal(vol, o.get_values())
        else:
            np.testing.assert_allclose(vol, o.get_values(), rtol=1e-5, atol=1e-6)


def test_shared_reciprocal_division_is_ieee_exact():
    """The kernels refine ONE reciprocal for p_x / p_z and p_y / p_z; on-device comparison against
    __fdiv_rn over 2^31 operand triples (random + near-power-of-two + repeated-mantissa cases)."""
    import ctypes
    from plant3dvision_b200 import _cabi
    bad = ctypes.c_uint64(12345)
    for seed in (0, 0x9E3779B9):
        _cabi.check(_cabi.lib().bp_selftest_division(0, 1 << 30, seed, ctypes.byref(bad)))
        assert bad.value == 0

"""GPU parity tests of the round-2 carve path: the block cull (summed-area tables of the first views, whole blocks
written off as carved, persistent dead flags, work list drained by persistent warps) and the two-tier masks (bit tiles
for the first views of a launch, the caller's raw uint8 masks read in place for the rest).  Everything goes through
the C ABI and is compared bit for bit with the CPU oracle and with the same library run with the cull switched off
(``BP_CARVE_CULL=0``: the round-1 kernel that sweeps every patch).
"""
import os

import numpy as np
import pytest

from conftest import split_cam

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def BP():
    from plant3dvision_b200.cl import Backprojection
    return Backprojection


@pytest.fixture(scope="module")
def Oracle():
    from oracle.backprojection_oracle import BackprojectionOracle
    return BackprojectionOracle


class env:
    """Temporarily set environment switches the library reads at every call."""

    def __init__(self, **kv):
        self.kv = {k: str(v) for k, v in kv.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *exc):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def synth_scene(name, scale, n_views=None):
    from plant3dvision_b200 import synth
    w = synth.workload(name, scale)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    views = [(*split_cam(w["cams"][v]), masks[v]) for v in range(len(w["cams"]))]
    return w, (views if n_views is None else views[:n_views])


def oracle_volume(Oracle, w, views, default_value=0, shape=None):
    ora = Oracle(shape or w["shape"], w["origin"], w["voxel_size"], type="carving", default_value=default_value)
    for v in views:
        ora.process_view(*v)
    return ora.get_values()


@pytest.mark.parametrize("name,scale", [("C5", 0.125), ("C2", 0.3), ("C4", 0.2)])
def test_cull_equals_no_cull_equals_oracle(BP, Oracle, name, scale):
    """Cull off / on with both passes (every packed view: BP_CULL_VIEWS=32) / on as a grid of this size gets it by
    default (the first pass only: few blocks) / on with 7 views - all equal to the oracle."""
    w, views = synth_scene(name, scale)
    want = oracle_volume(Oracle, w, views)
    for cull, n_views in ((1, 32), (1, None), (1, 7), (0, None)):
        kv = {"BP_CARVE_CULL": cull}
        if n_views:
            kv["BP_CULL_VIEWS"] = n_views
        with env(**kv):
            bp = BP(w["shape"], w["origin"], w["voxel_size"])
            bp.process_views(views)
            got = bp.get_values().copy()
            _, blocks, dead = bp.work_stats()
            bp.close()
        if cull and n_views == 32:  # (C4's cylinder fills a third of its box and the box fills the pictures: few blocks can go)
            assert blocks > 0 and dead > (0.3 if name == "C5" else 0.02) * blocks, f"the cull wrote off only {dead} of {blocks} blocks"
        elif not cull:
            assert dead == 0
        assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ from the oracle (cull {cull}, views {n_views})"


@pytest.mark.parametrize("shape", [[9, 9, 33], [8, 8, 32], [17, 7, 95], [1, 1, 1], [3, 70, 31], [40, 12, 200]])
def test_cull_block_edges(BP, Oracle, shape):
    """Grids that are not whole blocks (8 x 8 x 32 voxels) along any axis, and smaller than one block."""
    w, views = synth_scene("C2", 0.2, n_views=20)
    vs = 150.0 / max(shape)
    ww = dict(w, voxel_size=vs)
    want = oracle_volume(Oracle, ww, views, shape=shape)
    bp = BP(shape, w["origin"], vs)
    bp.process_views(views)
    got = bp.get_values().copy()
    bp.close()
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("p_fg", [1e-4, 3e-3])
def test_isolated_foreground_pixels(BP, Oracle, p_fg):
    """Masks that are background except for isolated single pixels: almost every tile is empty, so the cull fires
    everywhere it may, and a block must survive exactly when one of its voxels lands on one of the lone pixels -
    including pixels in the first / last row and column of a 32 x 32 tile."""
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 0.25)
    rng = np.random.default_rng(11)
    views = []
    for v in range(0, len(w["cams"]), 3):
        m = np.where(rng.random((w["H"], w["W"])) < p_fg, 255, 0).astype(np.uint8)
        m[31::32, 31::64] = 255   # tile corners
        m[32::64, 0::32] = 255
        views.append((*split_cam(w["cams"][v]), m))
    want = oracle_volume(Oracle, w, views)
    bp = BP(w["shape"], w["origin"], w["voxel_size"])
    bp.process_views(views)
    got = bp.get_values().copy()
    bp.close()
    assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ"
    # one view alone leaves survivors (voxels that land exactly on the lone pixels): the cull must keep their blocks
    bp = BP(w["shape"], w["origin"], w["voxel_size"])
    bp.process_views(views[:1])
    got1 = bp.get_values().copy()
    bp.close()
    want1 = oracle_volume(Oracle, w, views[:1])
    assert np.array_equal(got1, want1) and (want1 == 1).any()


def test_dead_blocks_persist_over_launches_and_clear_resets_them(BP, Oracle):
    """Host masks in several staging chunks (one carve launch per chunk) and the per-view streaming API: blocks written
    off by an earlier launch are skipped by the later ones; ``clear()`` brings every block back."""
    w, views = synth_scene("C5", 0.1, n_views=60)
    want = oracle_volume(Oracle, w, views)
    one_mask = views[0][3].nbytes
    with env(BP_STAGE_BYTES=3 * one_mask):
        bp = BP(w["shape"], w["origin"], w["voxel_size"])
        bp.process_views(views)
        np.testing.assert_array_equal(bp.get_values(), want)
        _, blocks, dead = bp.work_stats()
        assert dead > 0
        # a different scene after clear(): nothing of the first one may linger
        bp.clear()
        assert bp.work_stats()[2] == 0
        inv = [(i, r, t, np.invert(m)) for (i, r, t, m) in views[:12]]
        bp.process_views(inv)
        np.testing.assert_array_equal(bp.get_values(), oracle_volume(Oracle, w, inv))
        bp.close()
    # streaming: one launch per view
    bp = BP(w["shape"], w["origin"], w["voxel_size"])
    for v in views[:25]:
        bp.process_view(*v)
    np.testing.assert_array_equal(bp.get_values(), oracle_volume(Oracle, w, views[:25]))
    assert bp.work_stats()[2] > 0
    bp.close()


@pytest.mark.parametrize("invert", [False, True])
@pytest.mark.parametrize("head", [0, 1, 5, 40])
def test_two_tier_masks_device_resident(BP, Oracle, head, invert):
    """Device-resident uint8 masks: the first ``BP_PACK_HEAD`` views are bit-packed, the rest are read raw
    (ld.global.nc.u8 in the dense / medium / sparse phases), plain and inverted readings.  Odd image widths included
    (raw reads need no alignment)."""
    import torch
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 0.25)
    H, W = w["H"], w["W"] - 3
    cams = w["cams"].copy()
    masks = synth.render_masks(cams, w["capsules"], H, W).numpy()
    if invert:
        masks = np.invert(masks)
    n = 45
    views = [(*split_cam(cams[v]), np.invert(masks[v]) if invert else masks[v]) for v in range(n)]
    want = oracle_volume(Oracle, w, views)
    md = torch.from_numpy(masks[:n].copy()).cuda()
    ptrs = [md[v].data_ptr() for v in range(n)]
    hw = np.tile(np.array([[H, W]], dtype=np.int32), (n, 1))
    for cull in (1, 0):
        with env(BP_PACK_HEAD=head, BP_CARVE_CULL=cull):
            bp = BP(w["shape"], w["origin"], w["voxel_size"])
            bp.process_device_views(cams[:n].astype(np.float32), ptrs, hw, invert=invert)
            got = bp.get_values().copy()
            bp.close()
        assert np.array_equal(got, want), f"head {head} cull {cull}: {(got != want).sum()} voxels differ"


def test_two_tier_speckle_walks_every_phase_on_raw_masks(BP, Oracle):
    """Salt-and-pepper masks (see test_gpu_parity) read RAW from the second view on: the live count of every warp patch
    decays through dense, medium and sparse (multi-view) phases on ld.global.nc.u8 reads."""
    import torch
    from plant3dvision_b200 import synth
    w = synth.workload("C2", 0.2)
    rng = np.random.default_rng(5)
    n = len(w["cams"])
    masks = np.where(rng.random((n, w["H"], w["W"])) < 0.04, 0, 255).astype(np.uint8)
    views = [(*split_cam(w["cams"][v]), masks[v]) for v in range(n)]
    want = oracle_volume(Oracle, w, views)
    md = torch.from_numpy(masks).cuda()
    hw = np.tile(np.array([[w["H"], w["W"]]], dtype=np.int32), (n, 1))
    with env(BP_PACK_HEAD=1):
        bp = BP(w["shape"], w["origin"], w["voxel_size"])
        bp.process_device_views(w["cams"].astype(np.float32), [md[v].data_ptr() for v in range(n)], hw)
        got = bp.get_values().copy()
        bp.close()
    assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ"


@pytest.mark.parametrize("split,grab,lpt", [(1, 1, 1), (4, 1, 1), (16, 1, 1), (16, 4, 0), (33, 2, 1), (1000, 1, 0)])
def test_view_segments_and_scheduling_switches(BP, Oracle, split, grab, lpt):
    """The views after the first ``BP_VIEW_SPLIT`` of a launch run as segments of the still-live patches that any warp
    may take (carve_segments_kernel: merged into the state by an atomic byte-wise minimum, carving being order-free);
    ``BP_ITEM_GRAB`` / ``BP_LPT`` change only who does which patch.  Bit-exact whatever the split, on a scene whose
    hull voxels meet every view and on speckle masks whose patches decay slowly through every class."""
    import torch
    from plant3dvision_b200 import synth
    w, views = synth_scene("C5", 0.125, n_views=70)
    want = oracle_volume(Oracle, w, views)
    w2 = synth.workload("C2", 0.2)
    rng = np.random.default_rng(11)
    n = len(w2["cams"])
    speckle = np.where(rng.random((n, w2["H"], w2["W"])) < 0.03, 0, 255).astype(np.uint8)
    views2 = [(*split_cam(w2["cams"][v]), speckle[v]) for v in range(n)]
    want2 = oracle_volume(Oracle, w2, views2)
    md = torch.from_numpy(speckle).cuda()
    hw = np.tile(np.array([[w2["H"], w2["W"]]], dtype=np.int32), (n, 1))
    with env(BP_VIEW_SPLIT=split, BP_VIEW_SPLIT_MIN_REST=1, BP_ITEM_GRAB=grab, BP_LPT=lpt):
        bp = BP(w["shape"], w["origin"], w["voxel_size"])
        bp.process_views(views)
        got = bp.get_values().copy()
        bp.close()
        bp = BP(w2["shape"], w2["origin"], w2["voxel_size"])
        bp.process_views(views2)
        got2 = bp.get_values().copy()
        bp.clear()
        bp.process_device_views(w2["cams"].astype(np.float32), [md[v].data_ptr() for v in range(n)], hw)
        got3 = bp.get_values().copy()
        bp.close()
    assert np.array_equal(got, want), f"{(got != want).sum()} voxels differ (scene)"
    assert np.array_equal(got2, want2), f"{(got2 != want2).sum()} voxels differ (speckle, host masks)"
    assert np.array_equal(got3, want2), f"{(got3 != want2).sum()} voxels differ (speckle, raw device masks)"


def test_view_segments_default_value_and_chunks(BP, Oracle):
    """Segments over several launches of one scan (host masks in chunks of 20 views, split 4) and with a non-zero
    default: a voxel no view sees must keep the 'unseen' coding through the atomic merge."""
    w, views = synth_scene("C5", 0.1, n_views=60)
    one_mask = views[0][3].nbytes
    for d in (0, -3):
        want = oracle_volume(Oracle, w, views, default_value=d)
        with env(BP_VIEW_SPLIT=4, BP_VIEW_SPLIT_MIN_REST=1, BP_STAGE_BYTES=20 * one_mask):
            bp = BP(w["shape"], w["origin"], w["voxel_size"], default_value=d)
            bp.process_views(views)
            np.testing.assert_array_equal(bp.get_values(), want, err_msg=f"default {d}")
            bp.close()


@pytest.mark.parametrize("shape", [(9, 17, 33), (16, 8, 47), (23, 31, 64), (8, 8, 16)])
def test_lazy_clear_leaves_no_stale_state(BP, Oracle, shape):
    """``clear`` does not write the state when a block cull will (the first pass of the next launch writes every block:
    zeros over the blocks it writes off, "unseen" over the others).  Whatever looks at the state first must see a
    cleared grid: a second scan after a first one, ``get_values`` / the raw state right after ``clear``, the kernel
    without a cull, the per-view path - on shapes that are not multiples of the block or of the row padding."""
    w, views = synth_scene("C5", 0.06, n_views=24)
    org, vs = w["origin"], w["voxel_size"] * w["shape"][0] / max(shape)
    for lazy in (1, 0):
        with env(BP_LAZY_CLEAR=lazy):
            bp = BP(list(shape), org, vs, default_value=-3)
            ora = Oracle(list(shape), org, vs, type="carving", default_value=-3)
            for v in views[:12]:
                ora.process_view(*v)
            first = ora.get_values().copy()
            bp.process_views(views[:12])
            np.testing.assert_array_equal(bp.get_values(), first)
            # right after clear: an untouched grid, through every reader
            bp.clear()
            assert (bp.get_values() == -3).all()
            bp.clear()
            st = bp.state_tensor()
            bp.sync()
            assert int((st != 2).sum().item()) == 0
            bp.clear()
            tail, lo, hi = bp.get_values_tail()
            assert lo == hi == -3 and (tail == -3).all()
            # a second scan must not see anything of the first (other views carve other voxels)
            ora2 = Oracle(list(shape), org, vs, type="carving", default_value=-3)
            for v in views[12:]:
                ora2.process_view(*v)
            second = ora2.get_values()
            bp.clear()
            bp.process_views(views[12:])
            np.testing.assert_array_equal(bp.get_values(), second, err_msg=f"fused, lazy {lazy}")
            bp.clear()
            for v in views[12:]:
                bp.process_view(*v)
            np.testing.assert_array_equal(bp.get_values(), second, err_msg=f"view by view, lazy {lazy}")
            bp.clear()
            with env(BP_CARVE_CULL=0):
                bp.process_views(views[12:])
            np.testing.assert_array_equal(bp.get_values(), second, err_msg=f"no cull, lazy {lazy}")
            bp.close()


def test_cull_on_striped_and_slab_handles(BP, Oracle):
    """Plane-sharded handles: a cull block of 8 LOCAL planes spans non-contiguous global planes under stripes narrower
    than a block; the bounds must still contain every voxel."""
    w, views = synth_scene("C5", 0.1, n_views=40)
    want = oracle_volume(Oracle, w, views)
    nx = w["shape"][0]
    for lg, world in ((0, 3), (1, 2), (3, 4)):
        got = np.full_like(want, 77)
        for rank in range(world):
            s = BP(w["shape"], w["origin"], w["voxel_size"], stripes=(lg, world, rank))
            s.process_views(views)
            if len(s.planes):
                got[s.planes] = s.get_values()
            s.close()
        np.testing.assert_array_equal(got, want, err_msg=f"stripes lg {lg} world {world}")
    s = BP(w["shape"], w["origin"], w["voxel_size"], i_range=(nx // 3, nx - 5))
    s.process_views(views)
    np.testing.assert_array_equal(s.get_values(), want[nx // 3: nx - 5])
    s.close()


def test_default_values_with_cull(BP, Oracle):
    w, views = synth_scene("C5", 0.08, n_views=30)
    for d in (0, 5, -1, -3):
        bp = BP(w["shape"], w["origin"], w["voxel_size"], default_value=d)
        bp.process_views(views)
        np.testing.assert_array_equal(bp.get_values(), oracle_volume(Oracle, w, views, default_value=d), err_msg=str(d))
        bp.close()


def test_counted_work_drops_with_the_cull(BP, monkeypatch):
    """BP_COUNT_WORK=1 (counting variant of the kernels): the cull must cut the projections evaluated."""
    w, views = synth_scene("C5", 0.125)
    monkeypatch.setenv("BP_COUNT_WORK", "1")
    work = {}
    for cull in (0, 1):
        with env(BP_CARVE_CULL=cull):
            bp = BP(w["shape"], w["origin"], w["voxel_size"])
            bp.process_views(views)
            bp.sync()
            work[cull] = bp.work_stats()[0]
            bp.close()
    nominal = int(np.prod(w["shape"])) * len(views)
    assert 0 < work[1] < work[0] <= nominal * 1.01  # (padding lanes may count a little)


# ---------------------------------------------------------------------------------------------
# view skipping of the averaging path (avg_skip_kernel): exact by construction, checked bit for bit
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("log", [False, True])
@pytest.mark.parametrize("invert", [False, True])
def test_averaging_view_skipping_is_bit_exact(BP, Oracle, log, invert):
    """Sparse probability maps (mostly one byte value): whole (block, view) pairs are replaced by `sum += table[B]`.
    Same float32 sums, in the same view order, as the oracle and as the same library with BP_AVG_SKIP=0 - for hard and
    soft maps, log on / off, plain and inverted readings, and the fused task tail."""
    from plant3dvision_b200 import synth
    w = synth.workload("C5", 0.128)  # 768 x 512 maps: 16-byte rows (the batched tile scan needs them)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"], soft_px=2.0).numpy()
    if invert:
        masks = np.invert(masks)
    views = [(*split_cam(w["cams"][v]), masks[v]) for v in range(0, len(w["cams"]), 2)]
    ora = Oracle(w["shape"], w["origin"], w["voxel_size"], type="averaging", log=log, default_value=0.25)
    for (i, r, t, m) in views:
        ora.process_view(i, r, t, np.invert(m) if invert else m)
    want = ora.get_values()
    got = {}
    for skip in (1, 0):
        with env(BP_AVG_SKIP=skip):
            bp = BP(w["shape"], w["origin"], w["voxel_size"], type="averaging", log=log, default_value=0.25)
            bp.process_views(views, invert=invert)
            got[skip] = bp.get_values().copy()
            evaluated, blocks, _ = bp.work_stats()
            bp.close()
        nominal = int(np.prod(w["shape"])) * len(views)
        if skip:
            assert blocks > 0 and 0 < evaluated < 0.9 * nominal, f"only {nominal - evaluated} of {nominal} voxel-views skipped"
        else:
            assert evaluated == nominal
    np.testing.assert_array_equal(got[1], want)
    np.testing.assert_array_equal(got[0], want)


@pytest.mark.parametrize("blk_k", [8, 16, 64])
def test_averaging_view_skipping_shapes_and_device_maps(BP, Oracle, blk_k, monkeypatch):
    """Grids that are not whole skip blocks (8 x 32 x blk_k voxels), stripes, device-resident maps, maps whose rows are
    not 16-byte multiples (no skipping there, same result)."""
    monkeypatch.setenv("BP_AVG_BLK_K", str(blk_k))
    import torch
    from plant3dvision_b200 import synth
    w = synth.workload("C5", 0.096)  # 576 x 384 maps
    for (H, W) in ((w["H"], w["W"]), (w["H"] - 3, w["W"] - 5)):
        masks = synth.render_masks(w["cams"], w["capsules"], H, W, soft_px=1.5).numpy()[::3]
        cams = w["cams"][::3]
        views = [(*split_cam(cams[v]), masks[v]) for v in range(len(cams))]
        for shape in (w["shape"], [w["shape"][0] - 3, 37, 70], [5, 33, 65]):
            ora = Oracle(shape, w["origin"], w["voxel_size"], type="averaging", log=True)
            for v in views:
                ora.process_view(*v)
            want = ora.get_values()
            bp = BP(shape, w["origin"], w["voxel_size"], type="averaging", log=True)
            bp.process_views(views)
            np.testing.assert_array_equal(bp.get_values(), want, err_msg=f"host maps {H}x{W} {shape}")
            bp.clear()
            md = torch.from_numpy(masks.copy()).cuda()
            hw = np.tile(np.array([[H, W]], dtype=np.int32), (len(cams), 1))
            bp.process_device_views(cams.astype(np.float32), [md[v].data_ptr() for v in range(len(cams))], hw)
            np.testing.assert_array_equal(bp.get_values(), want, err_msg=f"device maps {H}x{W} {shape}")
            bp.close()
    got = np.zeros_like(want)
    for rank in range(3):
        s = BP(shape, w["origin"], w["voxel_size"], type="averaging", log=True, stripes=(1, 3, rank))
        s.process_views(views)
        if len(s.planes):
            got[s.planes] = s.get_values()
        s.close()
    np.testing.assert_array_equal(got, want)

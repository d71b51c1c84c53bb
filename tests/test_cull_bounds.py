"""bp_box_bounds_host (host only; csrc/bp_geom.h, the very function cull_blocks_kernel evaluates on the device): the
pixel bounds of a block of voxels must be CONSERVATIVE under the reference's own float32 projection
(kernels/backprojection.c:3-34).  The block cull of the carve path writes a whole block off as carved when its bounds
lie inside the picture and the tiles under them hold no foreground pixel; that is the reference's result only if
  (1) bounds valid and inside the picture  =>  EVERY voxel of the block hits the picture, and
  (2) every voxel that hits lands inside the bounds.
Checked against the oracle's per-voxel projection, plus a numpy emulation of the whole cull (summed-area table of
"tile has a foreground pixel") against the oracle's carve: no block the cull kills may hold a voxel the oracle keeps.
"""
import ctypes

import numpy as np
import pytest

from oracle import backprojection_oracle as bo
from plant3dvision_b200 import _cabi, synth

_F = ctypes.POINTER(ctypes.c_float)
_I = ctypes.POINTER(ctypes.c_int32)
_B = ctypes.POINTER(ctypes.c_uint8)
BLK = (8, 8, 32)


def box_bounds(box, origin, vs, cam, H, W):
    b = (ctypes.c_int32 * 6)(*[int(x) for x in box])
    o = np.ascontiguousarray(origin, dtype=np.float32)
    c = np.ascontiguousarray(cam, dtype=np.float32)
    lo, hi, ok = (ctypes.c_double * 2)(), (ctypes.c_double * 2)(), ctypes.c_int32()
    _cabi.check(_cabi.lib().bp_box_bounds_host(b, o.ctypes.data_as(_F), float(np.float32(vs)), c.ctypes.data_as(_F), H, W,
                                               lo, hi, ctypes.byref(ok)))
    return bool(ok.value), (lo[0], lo[1]), (hi[0], hi[1])


def project_all(shape, origin, vs, cam, H, W, i_offset=0):
    lib = bo._lib("restatement")
    n = int(np.prod(shape))
    px, py, hit = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.uint8)
    c = np.ascontiguousarray(cam, dtype=np.float32)
    o = np.ascontiguousarray(origin, dtype=np.float32)
    shp = np.ascontiguousarray(shape, dtype=np.int32)
    lib.oracle_project_view(H, W, px.ctypes.data_as(_I), py.ctypes.data_as(_I), hit.ctypes.data_as(_B),
                            c[0:4].ctypes.data_as(_F), c[4:13].ctypes.data_as(_F), c[13:16].ctypes.data_as(_F),
                            o.ctypes.data_as(_F), ctypes.c_float(float(np.float32(vs))), shp.ctypes.data_as(_I), int(i_offset))
    return px.reshape(shape), py.reshape(shape), hit.reshape(shape).astype(bool)


def blocks_of(shape):
    for bi in range(0, shape[0], BLK[0]):
        for bj in range(0, shape[1], BLK[1]):
            for bk in range(0, shape[2], BLK[2]):
                yield (bi, min(bi + BLK[0], shape[0]) - 1, bj, min(bj + BLK[1], shape[1]) - 1,
                       bk, min(bk + BLK[2], shape[2]) - 1)


def check_view(shape, origin, vs, cam, H, W):
    px, py, hit = project_all(shape, origin, vs, cam, H, W)
    n_valid = n_inside = 0
    for box in blocks_of(shape):
        ok, lo, hi = box_bounds(box, origin, vs, cam, H, W)
        if not ok:
            continue
        n_valid += 1
        sl = (slice(box[0], box[1] + 1), slice(box[2], box[3] + 1), slice(box[4], box[5] + 1))
        h = hit[sl]
        if h.any():
            x, y = px[sl][h], py[sl][h]
            assert lo[0] <= x.min() and x.max() <= hi[0] and lo[1] <= y.min() and y.max() <= hi[1], \
                f"box {box}: pixels [{x.min()}, {x.max()}] x [{y.min()}, {y.max()}] outside bounds {lo} {hi}"
        if lo[0] >= 0 and lo[1] >= 0 and hi[0] <= W - 1 and hi[1] <= H - 1:
            n_inside += 1
            assert h.all(), f"box {box}: bounds inside the picture but {(~h).sum()} voxels miss it"
    return n_valid, n_inside


@pytest.mark.parametrize("name,scale", [("C2", 0.15), ("C4", 0.12), ("C5", 0.06)])
def test_block_bounds_on_the_bench_workloads(name, scale):
    w = synth.workload(name, scale=scale)
    cams = np.asarray(w["cams"], dtype=np.float32)
    tot_valid = tot_inside = 0
    for v in range(0, len(cams), max(1, len(cams) // 6)):
        a, b = check_view(w["shape"], w["origin"], w["voxel_size"], cams[v], w["H"], w["W"])
        tot_valid += a
        tot_inside += b
    assert tot_valid > 0 and tot_inside > 0  # the cull has something to work with on these rigs


def test_block_bounds_for_random_cameras():
    rng = np.random.default_rng(21)
    H, W = 190, 300
    n_inside = 0
    for trial in range(40):
        shape = [int(rng.integers(3, 30)), int(rng.integers(3, 30)), int(rng.integers(3, 80))]
        vs = float(rng.uniform(0.2, 2.0))
        origin = (rng.uniform(-30, 30, 3) * (1000.0 if trial % 10 == 9 else 1.0)).astype(np.float32)
        centre = origin + 0.5 * vs * np.asarray(shape)
        dist = float(rng.choice([0.3, 1.0, 2.0, 5.0, 12.0])) * vs * max(shape)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        pos = centre + dist * d
        look = centre + rng.normal(size=3) * vs * max(shape) * 0.4 - pos
        zc = look / np.linalg.norm(look)
        xc = np.cross(rng.normal(size=3), zc)
        xc /= np.linalg.norm(xc)
        yc = np.cross(zc, xc)
        R = np.stack([xc, yc, zc])
        t = -R @ pos
        f = float(rng.uniform(60, 700))
        cam = np.concatenate([[f, f * rng.uniform(0.9, 1.1), W / 2 + rng.uniform(-40, 40), H / 2 + rng.uniform(-40, 40)],
                              R.reshape(-1), t]).astype(np.float32)
        n_inside += check_view(shape, origin, vs, cam, H, W)[1]
    assert n_inside > 50


def emulate_cull(w, views, n_cull):
    """numpy emulation of cull_blocks_kernel: dead[block] = some view among the first n_cull has the block's bounds
    inside the picture with no foreground tile under them."""
    shape = w["shape"]
    dead = {}
    sats = []
    for (cam, m) in views[:n_cull]:
        H, W = m.shape
        Wt, Ht = (W + 31) // 32, (H + 31) // 32
        c = np.zeros((Ht * 32, Wt * 32), dtype=bool)
        c[:H, :W] = m != 0
        anyt = c.reshape(Ht, 32, Wt, 32).any(axis=(1, 3))
        S = np.zeros((Ht + 1, Wt + 1), dtype=np.int64)
        S[1:, 1:] = anyt.cumsum(0).cumsum(1)
        sats.append(S)
    for box in blocks_of(shape):
        dead[box] = False
        for (cam, m), S in zip(views[:n_cull], sats):
            H, W = m.shape
            ok, lo, hi = box_bounds(box, w["origin"], w["voxel_size"], cam, H, W)
            if not ok or lo[0] < 0 or lo[1] < 0 or hi[0] > W - 1 or hi[1] > H - 1:
                continue
            tx0, tx1, ty0, ty1 = int(lo[0]) >> 5, int(hi[0]) >> 5, int(lo[1]) >> 5, int(hi[1]) >> 5
            if S[ty1 + 1, tx1 + 1] - S[ty0, tx1 + 1] - S[ty1 + 1, tx0] + S[ty0, tx0] == 0:
                dead[box] = True
                break
    return dead


@pytest.mark.parametrize("name,scale", [("C5", 0.07), ("C2", 0.15)])
def test_emulated_cull_never_kills_a_voxel_the_oracle_keeps(name, scale):
    w = synth.workload(name, scale=scale)
    cams = np.asarray(w["cams"], dtype=np.float32)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    sel = list(range(0, len(cams), 4))
    views = [(cams[v], masks[v]) for v in sel]
    ora = bo.BackprojectionOracle(w["shape"], w["origin"], w["voxel_size"], type="carving")
    for cam, m in views:
        ora.process_view(cam[0:4], cam[4:13], cam[13:16], m)
    vol = ora.get_values()
    dead = emulate_cull(w, views, n_cull=len(views))
    n_dead = 0
    for box, is_dead in dead.items():
        if is_dead:
            n_dead += 1
            blockvol = vol[box[0]:box[1] + 1, box[2]:box[3] + 1, box[4]:box[5] + 1]
            assert (blockvol == -1).all(), f"block {box} culled but the oracle keeps {(blockvol != -1).sum()} voxels"
    assert n_dead > (0.2 * len(dead) if name == "C5" else 0)

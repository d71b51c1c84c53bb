#!/usr/bin/env python3
"""CPU model of the block-level cull planned for carve_kernel: no GPU needed.

For the C5 workload at a reduced scale, every block of Bi x Bj x Bk voxels is tested against the views in the
kernel's farthest-point order: project its 8 corners, take the pixel bounding box widened by a margin, round it out to
whole 32 x 32 tiles, and ask a summed-area table of "tile has a foreground bit" whether the box is empty while lying
entirely inside the picture.  The first view that says so carves the whole block.  Reports
  * the fraction of blocks / voxels that survive the cull when it may use the first K views,
  * that no culled block holds a voxel the per-voxel oracle keeps (conservativeness),
  * the modelled warp instructions of the per-voxel kernel restricted to the surviving blocks, next to today's.
Usage: cull_model.py [scale] [Bi Bj Bk]      (0.25 -> 256^3 grid, 1500 x 1000 masks)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

PROLOGUE, DENSE, MEDIUM, SPARSE, DEAL = 102, 206, 121, 90, 70
CULL_TEST = 70  # thread instructions per block-view test (8 corner projections in float64 would be more; see kernel)


def death_positions(w, masks, cams, order):
    nx, ny, nz = w["shape"]
    vs, (ox, oy, oz) = np.float32(w["voxel_size"]), [np.float32(o) for o in w["origin"]]
    n_vox = nx * ny * nz
    death = np.full(n_vox, len(cams), dtype=np.int16)
    live = np.arange(n_vox, dtype=np.int64)
    for pos, v in enumerate(order):
        if live.size == 0:
            break
        i, r = np.divmod(live, ny * nz)
        j, k = np.divmod(r, nz)
        x, y, z = ox + i.astype(np.float32) * vs, oy + j.astype(np.float32) * vs, oz + k.astype(np.float32) * vs
        c = cams[v]
        R, t = c[4:13].reshape(3, 3), c[13:16]
        px = R[0, 0] * x + R[0, 1] * y + R[0, 2] * z + t[0]
        py = R[1, 0] * x + R[1, 1] * y + R[1, 2] * z + t[1]
        pz = R[2, 0] * x + R[2, 1] * y + R[2, 2] * z + t[2]
        with np.errstate(all="ignore"):
            u, vv = px / pz * c[0] + c[2], py / pz * c[1] + c[3]
        ok = (pz >= 0) & (u > -1) & (u < w["W"]) & (vv > -1) & (vv < w["H"])
        xi, yi = np.clip(u.astype(np.int64), 0, w["W"] - 1), np.clip(vv.astype(np.int64), 0, w["H"] - 1)
        dead = ok & (masks[v][yi, xi] == 0)
        death[live[dead]] = pos
        live = live[~dead]
    return death.reshape(nx, ny, nz)


def tile_sat(mask):
    H, W = mask.shape
    Wt, Ht = (W + 31) // 32, (H + 31) // 32
    coarse = np.zeros((Ht * 32, Wt * 32), dtype=bool)
    coarse[:H, :W] = mask != 0
    coarse = coarse.reshape(Ht, 32, Wt, 32).any(axis=(1, 3))
    sat = np.zeros((Ht + 1, Wt + 1), dtype=np.int64)
    sat[1:, 1:] = coarse.cumsum(0).cumsum(1)
    return sat, Wt, Ht


def block_cull_view(w, cam, sat, Wt, Ht, B, nb):
    """bool [nbi, nbj, nbk]: blocks this view carves entirely."""
    nx, ny, nz = w["shape"]
    vs = float(np.float32(w["voxel_size"]))
    ox, oy, oz = [float(np.float32(o)) for o in w["origin"]]
    c = cam.astype(np.float64)
    R, t = c[4:13].reshape(3, 3), c[13:16]
    lo = [np.arange(nb[a]) * B[a] for a in range(3)]
    hi = [np.minimum(lo[a] + B[a] - 1, (nx, ny, nz)[a] - 1) for a in range(3)]
    umin = np.full(nb, np.inf); umax = np.full(nb, -np.inf)
    vmin = np.full(nb, np.inf); vmax = np.full(nb, -np.inf)
    zmin = np.full(nb, np.inf)
    mag = np.zeros((3,) + tuple(nb))
    for ci in range(8):
        ii = (hi[0] if ci & 4 else lo[0])[:, None, None]
        jj = (hi[1] if ci & 2 else lo[1])[None, :, None]
        kk = (hi[2] if ci & 1 else lo[2])[None, None, :]
        x, y, z = ox + ii * vs, oy + jj * vs, oz + kk * vs
        wm = (abs(ox) + ii * vs, abs(oy) + jj * vs, abs(oz) + kk * vs)
        p = [R[a, 0] * x + R[a, 1] * y + R[a, 2] * z + t[a] for a in range(3)]
        for a in range(3):
            mag[a] = np.maximum(mag[a], abs(R[a, 0]) * wm[0] + abs(R[a, 1]) * wm[1] + abs(R[a, 2]) * wm[2] + abs(t[a]))
        with np.errstate(all="ignore"):
            u = p[0] / p[2] * c[0] + c[2]
            v = p[1] / p[2] * c[1] + c[3]
        umin, umax = np.minimum(umin, u), np.maximum(umax, u)
        vmin, vmax = np.minimum(vmin, v), np.maximum(vmax, v)
        zmin = np.minimum(zmin, p[2] - 1e-2 * mag[2])
    eps = 1.1920929e-7
    ok = zmin > 0
    with np.errstate(all="ignore"):
        pz_min = zmin + 1e-2 * mag[2]
        mu = 2.0 + 16 * eps * (abs(c[0]) * mag[0] + (w["W"] + abs(c[2])) * mag[2]) / pz_min + 8 * eps * (w["W"] + 2 * abs(c[2]))
        mv = 2.0 + 16 * eps * (abs(c[1]) * mag[1] + (w["H"] + abs(c[3])) * mag[2]) / pz_min + 8 * eps * (w["H"] + 2 * abs(c[3]))
    u0, u1 = np.floor(umin - mu), np.floor(umax + mu)
    v0, v1 = np.floor(vmin - mv), np.floor(vmax + mv)
    ok &= np.isfinite(u0) & np.isfinite(u1) & np.isfinite(v0) & np.isfinite(v1)
    ok &= (u0 >= 0) & (v0 >= 0) & (u1 <= w["W"] - 1) & (v1 <= w["H"] - 1)
    tx0, tx1 = np.clip(np.nan_to_num(u0) // 32, 0, Wt - 1).astype(np.int64), np.clip(np.nan_to_num(u1) // 32, 0, Wt - 1).astype(np.int64)
    ty0, ty1 = np.clip(np.nan_to_num(v0) // 32, 0, Ht - 1).astype(np.int64), np.clip(np.nan_to_num(v1) // 32, 0, Ht - 1).astype(np.int64)
    cnt = sat[ty1 + 1, tx1 + 1] - sat[ty0, tx1 + 1] - sat[ty1 + 1, tx0] + sat[ty0, tx0]
    return ok & (cnt == 0)


def patch_costs(death, nv):
    """Modelled warp instructions of every 4 x 32 patch (one plane) under the dense / medium / multi-view sparse phases
    (coarse: sparse priced at one pass per remaining view / g with the survivors at entry)."""
    nx, ny, nz = death.shape
    nyp, nzp = (ny + 3) // 4 * 4, (nz + 31) // 32 * 32
    d = np.full((nx, nyp, nzp), -1, dtype=np.int16)
    d[:, :ny, :nz] = death
    pat = d.reshape(nx, nyp // 4, 4, nzp // 32, 32).transpose(0, 1, 3, 2, 4).reshape(-1, 128)
    pat = np.sort(pat, axis=1)
    last = pat[:, -1].astype(np.int64)
    t_end = np.minimum(last + 1, nv)
    t_dense = np.minimum(np.maximum(pat[:, 63].astype(np.int64) + 1, 1), t_end)  # until <= 64 live
    t_med = np.minimum(np.maximum(pat[:, 95].astype(np.int64) + 1, t_dense), t_end)  # until <= 32 live
    t_dense = np.where(last < 0, 0, t_dense)
    t_med = np.where(last < 0, 0, t_med)
    rem = np.maximum(t_end - t_med, 0)
    surv = (pat[:, 96:].astype(np.int64) >= t_med[:, None]).sum(axis=1)
    g = np.where(surv > 0, 2 ** np.floor(np.log2(32 / np.maximum(surv, 1))), 1)
    sparse_passes = np.ceil(rem / g)
    cost = PROLOGUE + DENSE * t_dense + MEDIUM * (t_med - t_dense) + SPARSE * sparse_passes + DEAL * (rem > 0)
    shp = (nx, nyp // 4, nzp // 32)
    parts = {"prologue": np.full(shp, float(PROLOGUE)), "dense_first": (DENSE * np.minimum(t_dense, 1)).reshape(shp),
             "dense_later": (DENSE * np.maximum(t_dense - 1, 0)).reshape(shp), "medium": (MEDIUM * (t_med - t_dense)).reshape(shp),
             "sparse": (SPARSE * sparse_passes + DEAL * (rem > 0)).reshape(shp),
             "never_dies": (t_end >= nv).reshape(shp).astype(np.float64), "necessary_proj": np.minimum(pat.astype(np.int64) + 1, nv).clip(0).sum(axis=1).reshape(shp).astype(np.float64)}
    return cost.reshape(shp), parts


def main():
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.25
    B = [int(a) for a in sys.argv[2:5]] if len(sys.argv) >= 5 else [8, 8, 32]
    from plant3dvision_b200 import synth
    from carve_model import spread_order
    w = synth.workload("C5", scale=scale)
    t0 = time.time()
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    cams = w["cams"].astype(np.float32)
    order = spread_order(cams.astype(np.float64))
    nv = len(cams)
    print(f"rendered in {time.time() - t0:.1f} s, foreground {np.mean(masks != 0):.4f}", file=sys.stderr)
    death = death_positions(w, masks, cams, order)
    print(f"death positions in {time.time() - t0:.1f} s", file=sys.stderr)
    nx, ny, nz = w["shape"]
    nb = [-(-nx // B[0]), -(-ny // B[1]), -(-nz // B[2])]
    # last death position inside each block (nv = some voxel survives everything)
    pad = np.full((nb[0] * B[0], nb[1] * B[1], nb[2] * B[2]), -1, dtype=np.int16)
    pad[:nx, :ny, :nz] = death
    blk_last = pad.reshape(nb[0], B[0], nb[1], B[1], nb[2], B[2]).max(axis=(1, 3, 5))
    alive = np.ones(nb, dtype=bool)
    cull_pos = np.full(nb, nv, dtype=np.int64)
    wrong = 0
    out = {"scale": scale, "grid": [nx, ny, nz], "block": B, "blocks": int(np.prod(nb))}
    marks = {1, 2, 4, 8, 16, 32, 64, nv}
    for pos, v in enumerate(order):
        sat, Wt, Ht = tile_sat(masks[v])
        c = block_cull_view(w, cams[v], sat, Wt, Ht, B, nb) & alive
        wrong += int((c & (blk_last > pos)).sum())
        cull_pos[c] = pos
        alive &= ~c
        if pos + 1 in marks:
            out[f"blocks_alive_after_{pos + 1}_views"] = float(alive.mean())
    out["cull_wrong_decisions"] = wrong
    out["blocks_the_voxel_oracle_empties"] = float((blk_last < nv).mean())
    out["mean_views_tested_per_block"] = float((np.minimum(cull_pos, nv - 1) + 1).mean())
    cost, parts = patch_costs(death, nv)
    out["voxels_never_carved"] = float((death >= nv).mean())
    out["necessary_projections_per_voxel"] = float(np.minimum(death.astype(np.int64) + 1, nv).mean())
    total_now = float(cost.sum())
    # patches in surviving blocks only (patch grid: plane x (j/4) x (k/32))
    pj, pk = B[1] // 4, max(B[2] // 32, 1)
    for K in (16, 32, nv):
        live_blk = cull_pos >= K  # survived the first K views
        live_pat = np.repeat(np.repeat(np.repeat(live_blk, B[0], axis=0), pj, axis=1), pk, axis=2)[:cost.shape[0], :cost.shape[1], :cost.shape[2]]
        tests = np.minimum(cull_pos, K - 1) + 1
        total_cull = float((cost * live_pat).sum()) + float(tests.sum()) * CULL_TEST / 32.0
        out[f"K{K}_blocks_alive"] = float(live_blk.mean())
        if K == nv:
            for name, arr in parts.items():
                out[f"share_all_{name}"] = float(arr.sum() / (total_now if name not in ("never_dies", "necessary_proj") else arr.size))
                out[f"share_live_{name}"] = float((arr * live_pat).sum() / (total_now if name not in ("never_dies", "necessary_proj") else arr.size))
        out[f"K{K}_model_instr_ratio"] = total_cull / total_now
    out["model_warp_instr_now"] = total_now
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

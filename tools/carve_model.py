#!/usr/bin/env python3
"""CPU model of where carve_kernel's instructions go on the C5 workload (no GPU needed): for every voxel the position,
in the kernel's farthest-point view order, of the view that carves it; from that, per warp patch (4 j x 32 k voxels at
one i) the number of DENSE views (until <= 32 voxels are live) and SPARSE views (until none is), priced with the
SASS loop counts of tools/sass_loops.py (prologue 102, dense 219 / view, sparse 88 / view).  Prints the split and what
two candidate changes would save: (a) culling whole patches against a coarse mask pyramid, (b) compacting survivors
across warps.  Usage: carve_model.py [scale]  (0.25 -> 256^3 grid, 1500 x 1000 masks)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PROLOGUE, DENSE, SPARSE, DEAL = 102, 219, 88, 70


def spread_order(cams):
    n = len(cams)
    ax = cams[:, 10:13] / np.linalg.norm(cams[:, 10:13], axis=1, keepdims=True)
    far = np.full(n, 2.0)
    used = np.zeros(n, bool)
    order, cur = [], 0
    for _ in range(n):
        order.append(cur)
        used[cur] = True
        d = 1.0 - np.abs(ax @ ax[cur])
        far = np.minimum(far, d)
        cand = np.where(used, -1.0, far)
        cur = int(np.argmax(cand))
    return order


def main():
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.25
    from plant3dvision_b200 import synth
    w = synth.workload("C5", scale=scale)
    t0 = time.time()
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).numpy()
    print(f"rendered {masks.shape} in {time.time() - t0:.1f} s, foreground {np.mean(masks != 0):.4f}", file=sys.stderr)
    cams = w["cams"].astype(np.float32)
    nx, ny, nz = w["shape"]
    vs, (ox, oy, oz) = np.float32(w["voxel_size"]), [np.float32(o) for o in w["origin"]]
    order = spread_order(cams.astype(np.float64))
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
        if pos < 8 or pos % 40 == 0:
            print(f"view {pos:3d} (#{v:3d}): live {live.size} ({live.size / n_vox:.4%})", file=sys.stderr)
    nyp, nzp = (ny + 3) // 4 * 4, (nz + 31) // 32 * 32
    d = np.full((nx, nyp, nzp), -1, dtype=np.int16)  # padding voxels are inert (never live)
    d[:, :ny, :nz] = death.reshape(nx, ny, nz)
    pat = d.reshape(nx, nyp // 4, 4, nzp // 32, 32).transpose(0, 1, 3, 2, 4).reshape(-1, 128)
    pat = np.sort(pat, axis=1)
    nv = len(cams)
    last = pat[:, -1].astype(np.int64)                 # death position of the last voxel (nv = never)
    t_end = np.minimum(last + 1, nv)                   # views the warp looks at
    # dense phase: until <= 32 live, at least one view; live after view p = #(death > p)
    t_dense = np.maximum(pat[:, 95].astype(np.int64) + 1, 1)
    t_dense = np.minimum(t_dense, t_end)
    t_dense = np.where(last < 0, 0, t_dense)           # all-padding patches: prologue only
    t_sparse = np.maximum(t_end - t_dense, 0)
    cost = PROLOGUE + DENSE * t_dense + SPARSE * t_sparse + DEAL * (t_sparse > 0)
    n_w = len(pat)
    # survivors x views actually needed in the sparse phase (one lane-slot each)
    surv_views = 0
    for col in range(96, 128):
        dd = pat[:, col].astype(np.int64)
        surv_views += np.sum(np.clip(np.minimum(dd + 1, nv) - t_dense, 0, None))
    # (a) culling: a patch whose 128 voxels all die at the same view p could be removed by ONE coarse test at p
    same = (pat[:, 0] == pat[:, -1]) & (last < nv) & (last >= 0)
    cull_cost = np.where(same, PROLOGUE + 45 * (last + 1), cost)  # 45 instr per coarse patch-vs-view test
    out = {
        "scale": scale, "grid": [nx, ny, nz], "warps": int(n_w),
        "model_instr_per_warp": float(cost.mean()),
        "share_prologue": float(PROLOGUE / cost.mean()),
        "share_dense": float((DENSE * t_dense).mean() / cost.mean()),
        "share_sparse": float((SPARSE * t_sparse + DEAL * (t_sparse > 0)).mean() / cost.mean()),
        "mean_dense_views": float(t_dense.mean()), "mean_sparse_views": float(t_sparse.mean()),
        "warps_with_sparse_phase": float((t_sparse > 0).mean()),
        "warps_alive_at_the_end": float((last >= nv).mean()),
        "voxels_never_carved": float((death >= nv).mean()),
        "sparse_lane_utilisation": float(surv_views / max(1, (32 * t_sparse).sum())),
        "a_culling_patches_that_die_in_one_view": float(same.mean()),
        "a_culling_model_instr_per_warp": float(cull_cost.mean()),
        "b_compaction_model_instr_per_warp": float((PROLOGUE + DENSE * t_dense).mean() + (SPARSE + 10) * surv_views / 32 / n_w
                                                   + DEAL * (t_sparse > 0).mean()),
    }

    # (c) multi-view sparse phase (implemented): with c survivors, g = 2^floor(log2(32 / c)) views per pass;
    # the warp deals again when the live count allows g to double.
    sp = t_sparse > 0
    surv = pat[sp][:, 96:].astype(np.int64)            # death positions of the <= 32 survivors (sorted, -1 = padding)
    t = t_dense[sp].copy()
    end = t_end[sp]
    passes = np.zeros(len(t), dtype=np.int64)
    deals = np.zeros(len(t), dtype=np.int64)
    for _ in range(7):
        alive = (surv >= t[:, None]).sum(axis=1)
        act = (alive > 0) & (t < end)
        if not act.any():
            break
        lgv = np.where(act, np.floor(np.log2(32 / np.maximum(alive, 1))).astype(np.int64), 0)
        gv = 1 << lgv
        thr = np.where(lgv < 5, 32 >> (lgv + 1), 0)  # deal again once alive <= thr
        # time at which alive <= thr: the death position of the (thr+1)-th largest survivor, + 1
        idx = np.clip(32 - thr - 1, 0, 31)
        t_next = np.minimum(surv[np.arange(len(t)), idx] + 1, end)
        t_next = np.maximum(t_next, t)
        n_pass = -(-(t_next - t) // gv)
        passes += np.where(act, n_pass, 0)
        deals += act
        t = np.where(act, t + n_pass * gv, t)
    mv_cost = (PROLOGUE + DENSE * t_dense).astype(np.float64)
    mv_sparse = np.zeros(n_w)
    mv_sparse[sp] = 92 * passes + DEAL * deals
    out["c_multiview_model_instr_per_warp"] = float((mv_cost + mv_sparse).mean())
    out["c_multiview_mean_passes"] = float(passes.sum() / n_w)
    out["survivors_at_sparse_entry_mean"] = float((surv >= t_dense[sp][:, None]).sum(axis=1).mean())
    out["survivors_final_mean_in_warps_alive_at_end"] = float((surv[end == nv] >= nv).sum(axis=1).mean()) if (end == nv).any() else 0.0

    # where the dense views go: views a warp spends with 97..128 / 65..96 / 33..64 voxels live (view 0 always counts
    # in the first bucket)
    t96 = np.minimum(np.maximum(pat[:, 31].astype(np.int64) + 1, 1), t_dense)
    t64 = np.minimum(np.maximum(pat[:, 63].astype(np.int64) + 1, t96), t_dense)
    out["dense_views_97_128_live"] = float(t96.mean())
    out["dense_views_65_96_live"] = float((t64 - t96).mean())
    out["dense_views_33_64_live"] = float((t_dense - t64).mean())
    out["dense_views_after_the_first"] = float((t_dense - np.minimum(t_dense, 1)).mean())
    full = (pat[:, 0] >= nv)
    out["warps_fully_inside_the_hull"] = float(full.mean())

    # (a') the conservative patch-versus-view test planned for round 2, evaluated for real on the first views: project
    # the 4 corners of the patch (it is planar: one plane i), take the pixel bounding box with a 1-pixel margin, and
    # look the covered 32 x 32 tiles up in a coarse "tile has a set bit" bitmap (here via a summed-area table).
    # Checks that the test never culls a patch with a surviving voxel, and how many patches it retires.
    def corners(v):
        c = cams[v]
        R, t = c[4:13].reshape(3, 3), c[13:16]
        ii = np.arange(nx, dtype=np.float32)[:, None, None]
        jj = (np.arange(nyp // 4, dtype=np.float32) * 4)[None, :, None]
        kk = (np.arange(nzp // 32, dtype=np.float32) * 32)[None, None, :]
        us, vs_, zs = [], [], []
        for dj in (0, 3):
            for dk in (0, 31):
                x = ox + ii * vs
                y = oy + np.minimum(jj + dj, ny - 1) * vs
                z = oz + np.minimum(kk + dk, nz - 1) * vs
                px = R[0, 0] * x + R[0, 1] * y + R[0, 2] * z + t[0]
                py = R[1, 0] * x + R[1, 1] * y + R[1, 2] * z + t[1]
                pz = R[2, 0] * x + R[2, 1] * y + R[2, 2] * z + t[2]
                with np.errstate(all="ignore"):
                    us.append(px / pz * c[0] + c[2])
                    vs_.append(py / pz * c[1] + c[3])
                zs.append(pz)
        return np.stack(us), np.stack(vs_), np.stack(zs)

    Wt, Ht = (w["W"] + 31) // 32, (w["H"] + 31) // 32
    alive_pat = np.ones(n_w, dtype=bool)
    first = d.reshape(nx, nyp // 4, 4, nzp // 32, 32).transpose(0, 1, 3, 2, 4).reshape(-1, 128)
    culled_total, wrong = 0, 0
    for pos in range(4):
        v = order[pos]
        m = masks[v] != 0
        coarse = np.zeros((Ht * 32, Wt * 32), dtype=bool)
        coarse[:w["H"], :w["W"]] = m
        coarse = coarse.reshape(Ht, 32, Wt, 32).any(axis=(1, 3))
        sat = np.zeros((Ht + 1, Wt + 1), dtype=np.int64)
        sat[1:, 1:] = coarse.cumsum(0).cumsum(1)
        u, vv, zz = corners(v)
        ok = np.isfinite(u).all(0) & np.isfinite(vv).all(0) & (zz.min(0) > 1e-3)
        u0, u1 = np.floor(u.min(0) - 1), np.floor(u.max(0) + 1)
        v0, v1 = np.floor(vv.min(0) - 1), np.floor(vv.max(0) + 1)
        ok &= (u0 >= 0) & (v0 >= 0) & (u1 <= w["W"] - 1) & (v1 <= w["H"] - 1)
        tx0, tx1 = np.clip(u0 // 32, 0, Wt - 1).astype(np.int64), np.clip(u1 // 32, 0, Wt - 1).astype(np.int64)
        ty0, ty1 = np.clip(v0 // 32, 0, Ht - 1).astype(np.int64), np.clip(v1 // 32, 0, Ht - 1).astype(np.int64)
        cnt = sat[ty1 + 1, tx1 + 1] - sat[ty0, tx1 + 1] - sat[ty1 + 1, tx0] + sat[ty0, tx0]
        cull = (ok & (cnt == 0)).reshape(-1) & alive_pat
        # every real voxel of a culled patch must have died at or before this view
        bad = cull & ((first > pos).any(axis=1))
        wrong += int(bad.sum())
        dies_here = alive_pat & (first.max(axis=1) <= pos)
        out[f"cull_view{pos}_culled"] = float(cull.sum() / n_w)
        out[f"cull_view{pos}_patches_dead_after_it"] = float(dies_here.sum() / n_w)
        culled_total += int(cull.sum())
        alive_pat &= ~dies_here
    out["cull_wrong_decisions"] = wrong
    print(json.dumps(out))


if __name__ == "__main__":
    main()

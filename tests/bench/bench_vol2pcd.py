#!/usr/bin/env python3
"""Measure the volume -> point-cloud handoff on the C2 grid (301 x 301 x 551, 60 views):
GPU pipeline (device-timed, volume already in HBM as the carving state) next to the reference's own
scipy pipeline (oracle/vol2pcd_oracle.py = proc3d.py:490-570) on the host cores.  Prints one JSON line.

    python tests/bench/bench_vol2pcd.py [--scale 1.0] [--level-set 1.0] [--no-cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--level-set", type=float, default=1.0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    from plant3dvision_b200 import proc3d, synth
    from plant3dvision_b200.cl import Backprojection
    w = synth.workload(a.workload, a.scale)
    masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"], device="cuda").cpu().numpy()
    cams = w["cams"].astype(np.float32)
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving")
    bp.process_views([(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], masks[v]) for v in range(len(cams))])
    origin = np.asarray(w["origin"], dtype=np.float64)
    for _ in range(2):
        out = proc3d.vol2pcd(bp, origin, w["voxel_size"], a.level_set)
    tot, edt, wall = [], [], []
    for _ in range(a.reps):
        t0 = time.perf_counter()
        out = proc3d.vol2pcd(bp, origin, w["voxel_size"], a.level_set)
        wall.append(time.perf_counter() - t0)
        tot.append(out.timing["total_ms"])
        edt.append(out.timing["edt_ms"])
    nvox = int(np.prod(w["shape"]))
    line = {"what": "vol2pcd", "workload": f"{w['name']} carving volume {w['shape']} ({nvox} voxels), level_set {a.level_set}",
            "points": len(out), "band_voxels": out.timing["band_voxels"], "gpu_launches": out.timing["launches"],
            "gpu_ms_device": float(np.median(tot)), "gpu_edt_ms": float(np.median(edt)),
            "gpu_ms_wall_incl_alloc_and_fetch": 1e3 * float(np.median(wall)),
            "voxels_per_s_device": nvox / (np.median(tot) * 1e-3),
            # dense int16 traffic of the three EDT passes: read + write 2 B per voxel per pass (neighbour reads hit L1/L2)
            "edt_algorithmic_GBps": 3 * 4 * nvox / (np.median(edt) * 1e-3) / 1e9}
    if not a.no_cpu:
        from oracle import vol2pcd_oracle as vo
        vol = bp.get_values().copy()
        t0 = time.perf_counter()
        p, n = vo.vol2pcd(vol, origin, w["voxel_size"], a.level_set)
        cpu_s = time.perf_counter() - t0
        line["cpu_reference_s"] = cpu_s
        line["cpu_kind"] = "reference pipeline on scipy (vectorised point loop; the reference's joblib loop is slower)"
        line["same_point_count"] = bool(len(p) == len(out))
        line["max_abs_point_diff"] = float(np.abs(p - out.points).max()) if len(p) == len(out) and len(p) else None
        line["bit_equal"] = bool(len(p) == len(out) and np.array_equal(p, out.points) and np.array_equal(n, out.normals))
        line["speedup_device"] = cpu_s / (np.median(tot) * 1e-3)
    bp.close()
    print(json.dumps(line))


if __name__ == "__main__":
    main()

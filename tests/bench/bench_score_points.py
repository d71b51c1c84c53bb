#!/usr/bin/env python3
"""Measure SegmentedPointCloud scoring on the virtual-plant cameras and label maps: device kernel
time, wall time of the C-ABI call (uploads included), and the oracle (vectorised numpy restatement of the
reference's Python loop) on the host cores.  Prints one JSON line."""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    from test_score_points import LABELS, vp_points, vp_views
    z = np.load(os.path.join(ROOT, "tests", "golden", "virtual_plant_inputs.npz"))
    bbox = {"x": z["bbox"][0].tolist(), "y": z["bbox"][1].tolist(), "z": z["bbox"][2].tolist()}
    chans = ["background", "flower", "fruit", "leaf", "pedicel", "stem"]
    vp = {"cams": z["cams"], "bbox": bbox, "maps": {c: z[f"map_{c}"] for c in chans}}
    pts = vp_points(vp, n, seed=0, spread=1.0)
    views = vp_views(vp)
    p3 = importlib.import_module("plant-3d-vision_b200.proc3d")
    p3.score_points(pts[:1000], views, LABELS)  # context + module load
    wall, dev = [], []
    for _ in range(5):
        t0 = time.perf_counter()
        gl, gs, ms = p3.score_points(pts, views, LABELS)
        wall.append((time.perf_counter() - t0) * 1e3)
        dev.append(ms)
    from oracle.vol2pcd_oracle import score_points as oracle_score
    t0 = time.perf_counter()
    ol, os_ = oracle_score(pts, views, LABELS)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    n_proj = n * sum(1 for v in views if v[0] in LABELS)
    print(json.dumps({"workload": f"{n} points x {len(views)} label maps (virtual plant, 896x896, {len(LABELS)} labels)",
                      "projections": n_proj, "kernel_ms": float(np.median(dev)), "call_wall_ms": float(np.median(wall)),
                      "kernel_Gproj_per_s": n_proj / np.median(dev) / 1e6,
                      "oracle_numpy_ms": cpu_ms, "equal": bool(np.array_equal(gs, os_) and np.array_equal(gl, ol))}))


if __name__ == "__main__":
    main()

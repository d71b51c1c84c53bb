#!/usr/bin/env python3
"""Measure the Voxels task tail on one averaging label volume of the virtual-plant fixture at
voxel_size 0.1 ([232, 239, 1199] = 66.5 M voxels, log=True): the reference's host tail (get_values -> float64 stack
-> np.unique sanity check -> np.exp -> clip; cl.py:249-254, tasks/cl.py:166-174) against the fused read-back
(bp_get_values_tail).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    from conftest import split_cam, vp_shape_origin
    from plant3dvision_b200.cl import Backprojection
    vs = float(sys.argv[1]) if len(sys.argv) > 1 else 0.1
    z = np.load(os.path.join(ROOT, "tests", "golden", "virtual_plant_inputs.npz"))
    bbox = {"x": z["bbox"][0].tolist(), "y": z["bbox"][1].tolist(), "z": z["bbox"][2].tolist()}
    shape, origin = vp_shape_origin(bbox, vs)
    bp = Backprojection(shape, origin, vs, type="averaging", log=True)
    maps = z["map_stem"]
    bp.process_views([(*split_cam(z["cams"][v]), maps[v]) for v in range(len(maps))])
    bp.sync()
    n = int(np.prod(shape))

    def host_tail():
        result = np.zeros((1, *shape))
        result[0, :] = bp.get_values()
        t1 = time.perf_counter()
        degenerate = len(np.unique(result)) == 1
        t2 = time.perf_counter()
        vol = np.exp(result)
        vol[vol > 1] = 1.0
        return vol, degenerate, t2 - t1

    def fused_tail():
        result = np.zeros((1, *shape))
        _, lo, hi = bp.get_values_tail(out=result[0], float64=True, exp_clip=True)
        return result, lo == hi

    fused_tail()
    tf = []
    for _ in range(3):
        t0 = time.perf_counter()
        got, deg_f = fused_tail()
        tf.append(time.perf_counter() - t0)
    t0 = time.perf_counter()
    want, deg_h, t_unique = host_tail()
    th = time.perf_counter() - t0
    err = float(np.max(np.abs(got - want) / np.maximum(want, 1e-300)))
    print(json.dumps({"workload": f"one averaging label volume {shape} = {n} voxels, log=True, 18 views",
                      "reference_host_tail_s": th, "of_which_np_unique_s": t_unique, "fused_tail_s": float(np.median(tf)),
                      "speedup": th / float(np.median(tf)), "max_rel_err": err,
                      "clip_sets_equal": bool(np.array_equal(got == 1.0, want == 1.0)),
                      "same_sanity_verdict": bool(deg_f == deg_h), "cores": os.cpu_count()}))


if __name__ == "__main__":
    main()

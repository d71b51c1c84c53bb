#!/usr/bin/env python3
"""Torch-free quick check (a few seconds on the GPU box): carve random salt-and-pepper masks with the virtual-plant
cameras and compare with the oracle bit for bit - the idea of tests/test_gpu_parity.py::test_speckle_masks_* in a form
that fits a very short gpurun call.  Prints one JSON line."""
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
    from oracle.backprojection_oracle import BackprojectionOracle
    z = np.load(os.path.join(ROOT, "tests", "golden", "virtual_plant_inputs.npz"))
    bbox = {"x": z["bbox"][0].tolist(), "y": z["bbox"][1].tolist(), "z": z["bbox"][2].tolist()}
    vs = 0.3
    shape, origin = vp_shape_origin(bbox, vs)
    out = {"shape": shape}
    t0 = time.time()
    for p_zero in (0.08, 0.3):
        rng = np.random.default_rng(int(p_zero * 100))
        views = [(*split_cam(z["cams"][v]), np.where(rng.random((896, 896)) < p_zero, 0, 255).astype(np.uint8))
                 for v in range(18)]
        bp = Backprojection(shape, origin, vs, type="carving")
        bp.process_views(views)
        got = bp.get_values().copy()
        bp.close()
        ora = BackprojectionOracle(shape, origin, vs, type="carving")
        for v in views:
            ora.process_view(*v)
        want = ora.get_values()
        out[f"p{p_zero}"] = {"equal": bool(np.array_equal(got, want)), "differ": int((got != want).sum()),
                             "live": float((want != -1).mean())}
    out["seconds"] = time.time() - t0
    print(json.dumps(out))


if __name__ == "__main__":
    main()

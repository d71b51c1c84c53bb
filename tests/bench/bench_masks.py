#!/usr/bin/env python3
"""Measure the mask producer on a C2-sized scan: 60 RGB images of 1440 x 1080, `linear` [0, 1, 0],
threshold 0.3, a 3-step dilation (configs/geom_pipe_real.toml [Masks]; the footprint sequence here is cross x 3
because skimage's disk table is not in this image).  Device kernels, wall time of the C-ABI call (uploads and
read-back of the masks included), and the oracle (numpy / scipy restatement of the reference's per-image host code)
on a sample of the images.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    from test_masks import CROSS, plantish
    from plant3dvision_b200 import proc2d
    from oracle import masks_oracle as mo
    n, H, W = 60, 1080, 1440
    rng = np.random.default_rng(0)
    base = plantish(rng, H, W)
    base[:, :, 1] = (base[:, :, 1].astype(np.int32) * 2 // 5 + (base[:, :, 1] > 90) * 100).astype(np.uint8)  # dim background
    imgs = [np.ascontiguousarray(np.roll(base, (int(rng.integers(0, H)), int(rng.integers(0, W))), axis=(0, 1))) for _ in range(n)]
    seq = ((CROSS, 3),)
    proc2d.masks(imgs[:2], "linear", [0, 1, 0], 0.3, seq)  # context, module load
    wall, dev = [], []
    for _ in range(3):
        t0 = time.perf_counter()
        got = proc2d.masks(imgs, "linear", [0, 1, 0], 0.3, seq)
        wall.append(time.perf_counter() - t0)
        dev.append(proc2d.masks.last_ms)
    t0 = time.perf_counter()
    ptr, *_rest, ms_dev = proc2d.masks(imgs, "linear", [0, 1, 0], 0.3, seq, keep_on_device=True)
    wall_dev = time.perf_counter() - t0
    proc2d.device_free(ptr)
    sample = 4
    t0 = time.perf_counter()
    want = [mo.masks_filter(im, "linear", [0, 1, 0], 0.3, seq) for im in imgs[:sample]]
    cpu = (time.perf_counter() - t0) / sample * n
    px = n * H * W
    print(json.dumps({"workload": f"{n} RGB images {W}x{H}, linear [0,1,0] > 0.3, dilation cross x 3",
                      "kernels_ms": float(np.median(dev)), "kernel_Gpx_per_s": px / np.median(dev) / 1e6,
                      "kernel_input_GB_per_s": px * 3 / np.median(dev) / 1e6,
                      "call_wall_ms": float(np.median(wall)) * 1e3, "call_wall_masks_left_on_device_ms": wall_dev * 1e3,
                      "oracle_numpy_scipy_ms_extrapolated": cpu * 1e3, "oracle_sample": f"{sample} of {n} images",
                      "equal": bool(all(np.array_equal(g, w) for g, w in zip(got, want))),
                      "foreground_fraction": float(np.mean([g.mean() / 255 for g in got]))}))


if __name__ == "__main__":
    main()

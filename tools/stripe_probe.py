#!/usr/bin/env python3
"""Per-rank cost of the device-resident C5 carving step at world sizes 1 / 2 / 4 / 8, measured on ONE GPU: rank 0's
striped handle alone (the other ranks do the same work on the same masks - stripes of 8 planes dealt round robin balance
to a few percent).  Sweeps the scheduling switches of the carve path (BP_ITEM_GRAB, BP_LPT, BP_VIEW_SPLIT; 1000 = no view segments).
Usage: python tools/stripe_probe.py [worlds, e.g. 1,2,8] > gpurun_out/stripe_probe.txt"""
import itertools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from plant3dvision_b200 import synth  # noqa: E402
from plant3dvision_b200.cl import Backprojection  # noqa: E402


def main():
    worlds = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "1,2,8").split(",")]
    w = synth.workload("C5", 1.0)
    n, H, W = len(w["cams"]), w["H"], w["W"]
    dev = torch.device("cuda", 0)
    masks = torch.empty((n, H, W), dtype=torch.uint8, device=dev)
    for a in range(0, n, 16):
        masks[a:a + 16] = synth.render_masks(w["cams"][a:a + 16], w["capsules"], H, W, device=dev)
    ptrs = [masks[v].data_ptr() for v in range(n)]
    cams = w["cams"].astype(np.float32)
    hw = np.tile(np.array([[H, W]], dtype=np.int32), (n, 1))
    combos = [dict(BP_ITEM_GRAB=g, BP_LPT=l, BP_LIVE_ORDER=1, BP_PACK_HEAD="auto", BP_VIEW_SPLIT=sp)
              for g, l, sp in ((1, 1, 16), (1, 1, 1000), (4, 1, 16), (2, 1, 16), (1, 0, 16), (1, 1, 8), (1, 1, 32))]
    if os.environ.get("PROBE_SPLITS"):  # short sweep: grab 1, LPT on, the given view splits
        combos = [dict(BP_ITEM_GRAB=int(gr), BP_LPT=1, BP_LIVE_ORDER=1, BP_PACK_HEAD=hd, BP_VIEW_SPLIT=int(sp))
                  for sp in os.environ["PROBE_SPLITS"].split(",") for hd in os.environ.get("PROBE_HEADS", "auto").split(",")
                  for gr in os.environ.get("PROBE_GRABS", "1").split(",")]
        if os.environ.get("PROBE_CULLS"):
            combos = [dict(c, BP_CULL_VIEWS=int(cv)) for c in combos for cv in os.environ["PROBE_CULLS"].split(",")]
    print("world  grab lpt split head | step_ms kernel_ms ingest_ms")
    for world in worlds:
        for c in combos:
            for k, v in c.items():
                if v == "auto":
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = str(v)
            bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], stripes=(3, world, 0)) if world > 1 else \
                Backprojection(w["shape"], w["origin"], w["voxel_size"])
            ext = torch.cuda.ExternalStream(bp.stream(), device=dev)
            for _ in range(3):
                bp.clear()
                bp.process_device_views(cams, ptrs, hw)
            bp.sync()
            bp.timing()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            steps = 5
            for _ in range(steps):
                bp.clear()
                bp.process_device_views(cams, ptrs, hw)
            e1.record(ext)
            bp.sync()
            torch.cuda.synchronize()
            k_ms, i_ms, _ = bp.timing()
            print((f"cull {c['BP_CULL_VIEWS']:2} " if 'BP_CULL_VIEWS' in c else "") + f"{world:5d}  {c['BP_ITEM_GRAB']:4} {c['BP_LPT']:3} {c['BP_VIEW_SPLIT']:5} {str(c['BP_PACK_HEAD']):>4} | "
                  f"{e0.elapsed_time(e1) / steps:7.3f} {k_ms / steps:9.3f} {i_ms / steps:9.3f}", flush=True)
            bp.close()


if __name__ == "__main__":
    main()

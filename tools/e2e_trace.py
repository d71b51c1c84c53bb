#!/usr/bin/env python3
"""Where does the N-GPU end-to-end step go?  Runs the bench's e2e carving step (C5, host masks, plane-sharded, sharded
ingest) a few times as timed by the bench, then once with a device synchronisation after every phase (un-overlapped
phase costs).  torchrun --nproc-per-node N tools/e2e_trace.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from plant3dvision_b200 import synth  # noqa: E402
from plant3dvision_b200.parallel import DistributedBackprojection  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if os.environ.get("BIND", "0") == "1":
        from plant3dvision_b200.parallel import bind_to_gpu_numa_node
        cpus = bind_to_gpu_numa_node(int(os.environ.get("LOCAL_RANK", 0)))
        print(f"rank {rank}: bound to {len(cpus) if cpus else None} CPUs {cpus[:4] if cpus else ''}..", flush=True)
    else:
        print(f"rank {rank}: affinity {len(os.sched_getaffinity(0))} CPUs", flush=True)
    dist.init_process_group("nccl")
    w = synth.workload("C5", 1.0)
    dev_masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"])
    host = torch.empty(dev_masks.shape, dtype=torch.uint8).pin_memory()
    host.copy_(dev_masks)
    del dev_masks
    masks = host.numpy()
    cams = w["cams"].astype(np.float32)
    views = [(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], masks[v]) for v in range(len(cams))]
    bp = DistributedBackprojection(w["shape"], w["origin"], w["voxel_size"], sharding="stripes", stripe_lg=3, device=int(os.environ.get("LOCAL_RANK", 0)))

    def step():
        bp.clear()
        bp.process_views(views)
        return bp.get_values(gather=False)

    for _ in range(2):
        step()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        step()
    dist.barrier()
    ms = 1e3 * (time.perf_counter() - t0) / 4
    if rank == 0:
        print(f"world {world}: e2e step {ms:.2f} ms (ingest {bp.last_ingest})", flush=True)
    # phase by phase
    bp.trace = []
    t_a = time.perf_counter()
    bp.clear()
    torch.cuda.synchronize()
    t_b = time.perf_counter()
    bp.process_views(views)
    torch.cuda.synchronize()
    t_c = time.perf_counter()
    bp.get_values(gather=False)
    t_d = time.perf_counter()
    for rr in range(world):
      dist.barrier()
      if rank == rr:
        print(f"rank {rank}:  clear {1e3 * (t_b - t_a):.2f} ms; process_views {1e3 * (t_c - t_b):.2f} ms; get_values {1e3 * (t_d - t_c):.2f} ms (serialised)")
        last = t_b
        for label, t in bp.trace:
            print(f"    {label:60s} {1e3 * (t - last):8.2f} ms", flush=True)
            last = t
    bp.trace = None
    # the two transfers alone
    x = torch.empty(len(masks[0::world]) * masks[0].size, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    x.view(-1, masks[0].size).copy_(host[0::world].view(-1, masks[0].size), non_blocking=True) if world == 1 else [x[q * masks[0].size:(q + 1) * masks[0].size].copy_(host[v].view(-1), non_blocking=True) for q, v in enumerate(range(rank, len(masks), world))]
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if rank == 0:
        print(f"  plain H2D of this rank's {x.numel() / 1e6:.0f} MB of masks, all ranks at once: {1e3 * (t1 - t0):.2f} ms = {x.numel() / (t1 - t0) / 1e9:.1f} GB/s")
    # the download alone: this rank's share of the int32 volume, all ranks at once, into pinned memory
    nloc = int(np.prod(bp.local.local_shape))
    vol_d = torch.zeros(nloc, dtype=torch.int32, device="cuda")
    vol_h = torch.empty(nloc, dtype=torch.int32).pin_memory()
    for _ in range(2):
        torch.cuda.synchronize(); dist.barrier()
        t0 = time.perf_counter()
        vol_h.copy_(vol_d, non_blocking=True)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
    if rank == 0:
        print(f"  plain D2H of this rank's {4 * nloc / 1e6:.0f} MB of int32 volume, all ranks at once: {1e3 * (t1 - t0):.2f} ms = {4 * nloc / (t1 - t0) / 1e9:.1f} GB/s per rank, {world * 4 * nloc / (t1 - t0) / 1e9:.1f} GB/s aggregate")
    # both directions at once
    s2 = torch.cuda.Stream()
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    vol_h.copy_(vol_d, non_blocking=True)
    with torch.cuda.stream(s2):
        [x[q * masks[0].size:(q + 1) * masks[0].size].copy_(host[v].view(-1), non_blocking=True) for q, v in enumerate(range(rank, len(masks), world))]
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if rank == 0:
        print(f"  both at once: {1e3 * (t1 - t0):.2f} ms for {x.numel() / 1e6:.0f} MB up + {4 * nloc / 1e6:.0f} MB down per rank")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""bench.py — voxel-views/s of the back-projection hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload C5|C2|C4] [--scale S] [--mode carving|averaging]

Default workload (N=1 and N>1): C5 of BASELINE.md §3 — carving a 1024^3 grid from 240 views of
6000x4000 masks (the configuration north_star's target is quoted on; 1 GiB of int8 state + 5.8 GB
of masks fit one B200).  With N ranks the grid is cut into N axis-0 slabs, every
rank processes all views for its slab, no collective on the data path: total work is fixed
("scaling": "strong").

One step = clear the grid + ingest (bit-pack) every mask + the fused carve kernel over all views.
  value : whole-job voxel-views/s with the uint8 masks already resident in HBM, device-timed
  e2e   : same step through the public Backprojection API from pinned HOST masks, including the
          H2D copies and the get_values() D2H of the int32 volume, wall-clock between syncs
  roofline / cpu_baseline / clocks / gpu_launches : see DESIGN.md §Measurement
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "voxel-views/sec"
UNIT = "voxel-views/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C5", choices=["C5", "C2", "C4"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink grid and images (debug only)")
    ap.add_argument("--mode", default="carving", choices=["carving", "averaging"])
    ap.add_argument("--cpu-planes", type=int, default=0, help="planes in the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-averaging", action="store_true", help="carving run: skip the secondary averaging leg")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(mode):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(mode)
    return None


def ncu_counters(mode):
    """The other ncu counters of the same committed capture (L2 bytes and GB/s, issue-slot utilisation, hit
    rates), or None.  Static evidence from profiles/, not measured by this run."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return (json.load(f).get("ncu") or {}).get(mode)
    except (OSError, ValueError):
        return None


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region, sampled every ~5 ms through NVML in a
    thread (same counters as the nvidia-smi clocks line of B200_PROFILING.md; nvidia-smi itself cannot
    sample a 100 ms region)."""

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples, self.reasons, self.power = [], set(), []
        self.max_mhz = None
        self._stop = False
        self._t = None
        self._nv = None

    def _run(self):
        nv, h = self._nv, self._h
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self._stop:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        try:
            import pynvml as nv
            import threading
            nv.nvmlInit()
            # NVML enumerates physical GPUs; map through CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = self.idx
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                phys = int(vis.split(",")[self.idx])
            self._h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM))
            self._nv = nv
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        except Exception:
            self._nv = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        if self._nv is None:
            return out
        self._stop = True
        self._t.join(timeout=2)
        if self.samples:
            out.update(sm_mhz=float(np.median(self.samples)), reasons=sorted(self.reasons), samples=len(self.samples),
                       power_w_max=max(self.power) if self.power else None)
        return out


def sample_planes(nx, n):
    """Evenly spread axis-0 planes used as the bounded CPU sample."""
    n = max(1, min(n, nx))
    return sorted({int((q + 0.5) * nx / n) for q in range(n)})


# ---------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/)
# ---------------------------------------------------------------------------------------------
def cpu_leg(w, masks_np, mode, planes, kind, gpu_planes=None):
    """Time the CPU implementation on `planes` (each run as a 1-plane slab, all views).
    kind == "port": oracle/bp_oracle.c restatement (global plane index, comparable bit for bit);
    kind == "reference": the reference's own kernel source from oracle/_ref (origin shifted to the
    plane), including the int32 widening process_view does per view (cl.py:215)."""
    import ctypes
    from oracle import backprojection_oracle as bo
    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    cams = w["cams"].astype(np.float32)
    ny, nz = w["shape"][1], w["shape"][2]
    n_views = len(cams)
    F = ctypes.POINTER(ctypes.c_float)
    I = ctypes.POINTER(ctypes.c_int32)
    B = ctypes.POINTER(ctypes.c_uint8)
    origin = np.array(w["origin"], dtype=np.float32)
    vs = np.float32(w["voxel_size"])
    shape1 = np.array([1, ny, nz], dtype=np.int32)
    lut = None
    if mode == "averaging":
        lut = np.ascontiguousarray(bo.img_as_float32(np.arange(256, dtype=np.uint8)), dtype=np.float32)
    lib = bo._lib("restatement" if kind == "port" else "ref_off")
    ok = True
    t_total = 0.0
    for pi, i0 in enumerate(planes):
        if mode == "carving":
            state = np.zeros((1, ny, nz), dtype=np.int32)
        else:
            state = np.zeros((1, ny, nz), dtype=np.float32)
        t0 = time.perf_counter()
        for v in range(n_views):
            m = masks_np[v]
            H, W = m.shape
            c = cams[v]
            if kind == "port":
                if mode == "carving":
                    lib.oracle_carve_view(m.ctypes.data_as(B), H, W, state.ctypes.data_as(I), c[0:4].ctypes.data_as(F),
                                          c[4:13].ctypes.data_as(F), c[13:16].ctypes.data_as(F),
                                          origin.ctypes.data_as(F), ctypes.c_float(float(vs)),
                                          shape1.ctypes.data_as(I), int(i0))
                else:
                    lib.oracle_average_view_lut(m.ctypes.data_as(B), lut.ctypes.data_as(F), H, W,
                                                state.ctypes.data_as(F), c[0:4].ctypes.data_as(F),
                                                c[4:13].ctypes.data_as(F), c[13:16].ctypes.data_as(F),
                                                origin.ctypes.data_as(F), ctypes.c_float(float(vs)),
                                                shape1.ctypes.data_as(I), int(i0))
            else:
                volinfo = np.array([origin[0] + np.float32(i0) * vs, origin[1], origin[2], vs], dtype=np.float32)
                if mode == "carving":
                    mh = np.ascontiguousarray(m, dtype=np.int32)  # cl.py:215
                    lib.ref_carve_view(mh.ctypes.data_as(I), H, W, state.ctypes.data_as(I), c[0:4].ctypes.data_as(F),
                                       c[4:13].ctypes.data_as(F), c[13:16].ctypes.data_as(F),
                                       volinfo.ctypes.data_as(F), shape1.ctypes.data_as(I))
                else:
                    mh = np.ascontiguousarray(bo.img_as_float32(m), dtype=np.float32)  # cl.py:205-215
                    lib.ref_average_view(mh.ctypes.data_as(F), H, W, state.ctypes.data_as(F),
                                         c[0:4].ctypes.data_as(F), c[4:13].ctypes.data_as(F),
                                         c[13:16].ctypes.data_as(F), volinfo.ctypes.data_as(F),
                                         shape1.ctypes.data_as(I))
        t_total += time.perf_counter() - t0
        if gpu_planes is not None and kind == "port":
            ok = ok and bool(np.array_equal(state[0], gpu_planes[pi]))
    vv = float(len(planes)) * ny * nz * n_views
    return vv / t_total, t_total, cores, ok


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from plant3dvision_b200 import synth
    from oracle import backprojection_oracle as bo
    import torch
    w = synth.workload(args.workload, args.scale)
    kind = "reference" if bo.have_backend("ref_off") else "port"
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    masks = render_all(w, dev, args.mode, pinned=False)
    # the reference kernel source (one pass per view over int32 / float32 images it first widens) runs ~20x
    # slower than the port: size the per-step sample so that `--steps 5 --warmup 3` ends within ~2 minutes
    n_pl = args.cpu_planes or (max(2, auto_planes(w, args.mode) // 5) if kind == "reference" else auto_planes(w, args.mode))
    planes = sample_planes(w["shape"][0], n_pl)
    for _ in range(min(args.warmup, 1)):
        cpu_leg(w, masks, args.mode, planes[:1], kind)
    times, vals = [], []
    for _ in range(args.steps):
        v, t, cores, _ = cpu_leg(w, masks, args.mode, planes, kind)
        vals.append(v)
        times.append(t)
    value = float(len(planes)) * w["shape"][1] * w["shape"][2] * len(w["cams"]) * len(times) / sum(times)
    sample = (f"{len(planes)} of {w['shape'][0]} axis-0 planes (evenly spread) x all {len(w['cams'])} views, "
              f"per step; OpenMP over {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 projection, i32 state" if args.mode == "carving" else "f32", "data": "synthetic",
        "config": workload_config(w, args, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


def auto_planes(w, mode):
    # ~10-30 s of CPU work: carving skips carved voxels early, averaging touches every voxel-view
    per_plane = w["shape"][1] * w["shape"][2] * len(w["cams"])
    budget = 6e9 if mode == "carving" else 1.5e9
    return int(max(1, min(16, budget // max(per_plane, 1))))


def workload_config(w, args, n_gpus):
    return {
        "workload": f"{w['name']} {args.mode}: grid {w['shape'][0]}x{w['shape'][1]}x{w['shape'][2]}, "
                    f"{len(w['cams'])} views of {w['W']}x{w['H']} uint8 masks (synthetic phantom silhouettes)",
        "grid": w["shape"], "views": len(w["cams"]), "mask_hw": [w["H"], w["W"]], "mode": args.mode,
        "sharding": (f"axis-0 stripes (blocks of 8 planes dealt round robin) x{n_gpus}; masks ingested 1/{n_gpus} per rank + all-gather of the packed tiles "
                     f"over NVLink (NCCL)") if n_gpus > 1 else "single GPU",
        "l2": "inputs larger than L2 (masks + grid >> 126 MB); grid re-seeded every step",
        "scale": args.scale,
    }


def render_all(w, dev, mode, pinned):
    """uint8 masks [n, H, W] in host memory (numpy); pinned when they feed the e2e leg."""
    import torch
    from plant3dvision_b200 import synth
    n, H, W = len(w["cams"]), w["H"], w["W"]
    host = torch.empty((n, H, W), dtype=torch.uint8, pin_memory=(pinned and torch.cuda.is_available()))
    chunk = 16
    soft = 2.0 if mode == "averaging" else 0.0
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        host[a:b].copy_(synth.render_masks(w["cams"][a:b], w["capsules"], H, W, device=dev, soft_px=soft))
    return host.numpy()


def measure_mode(args, mode, w, cams, masks_np, ptrs, hw, views_host, world, rank, local, dev, with_cpu=True):
    """Device-resident throughput, roofline, e2e and CPU sample of one mode on the already rendered masks."""
    import torch
    import torch.distributed as dist
    from plant3dvision_b200.cl import Backprojection
    nx, ny, nz = w["shape"]
    n_views = len(w["cams"])
    H, W = w["H"], w["W"]
    i0, i1 = rank * nx // world, (rank + 1) * nx // world
    if world > 1:
        # axis-0 slabs; mask ingest sharded over the ranks + one all-gather over NVLink (parallel.py)
        from plant3dvision_b200.parallel import DistributedBackprojection
        bp = DistributedBackprojection(w["shape"], w["origin"], w["voxel_size"], type=mode, log=False,
                                       sharding="stripes", device=local,
                                       stripe_lg=int(os.environ.get("BP_STRIPE_LG", "3")))
        eng = bp.local
    else:
        bp = eng = Backprojection(w["shape"], w["origin"], w["voxel_size"], type=mode, log=False, device=local,
                                  i_range=(i0, i1))
    ext = torch.cuda.ExternalStream(eng.stream(), device=dev)

    def step_resident():
        bp.clear()
        bp.process_device_views(cams, ptrs, hw)

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_resident()
    barrier()
    eng.timing()  # reset the library's event accumulators
    _, _, launches0 = eng.timing()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(args.steps):
        step_resident()
    e1.record(ext)
    barrier()
    ms_total = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    kernel_ms, ingest_ms, launches1 = eng.timing()
    t = torch.tensor([ms_total, kernel_ms, ingest_ms, float(launches1 - launches0)], dtype=torch.float64, device=dev)
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    ms_step = float(tmax[0]) / args.steps
    total_vv = float(nx) * ny * nz * n_views
    value = total_vv / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel (carve / average), SURVEY §8(d) algorithmic bytes
    peak, peak_src = peaks()
    bytes_per_vv = 4.0 if mode == "carving" else 8.0
    mask_bytes = n_views * (H * W / 8.0 if mode == "carving" else float(H * W))
    n_local = len(eng.planes)  # axis-0 planes this rank holds
    algo_bytes = bytes_per_vv * float(n_local) * ny * nz * n_views + mask_bytes
    k_ms = float(tmax[1]) / args.steps  # one launch per step (n_views <= 256)
    n_launch_per_step = (n_views + 255) // 256
    achieved = algo_bytes / (k_ms / n_launch_per_step * 1e-3) / 1e9 / n_launch_per_step
    fused_floor = (1.0 if mode == "carving" else 8.0) * float(n_local) * ny * nz + mask_bytes
    roofline = {
        "bound": "hbm", "kernel": "carve_kernel" if mode == "carving" else "average_kernel<u8>",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": ncu_traffic(mode), "ncu": ncu_counters(mode), "peak_source": peak_src,
        "kernel_ms_per_launch": k_ms / n_launch_per_step, "ingest_ms_per_step": float(tmax[2]) / args.steps,
        "algorithmic_bytes_per_launch": algo_bytes,
        "note": "algorithmic bytes = SURVEY 8(d) per-view streaming traffic (4 B/voxel-view carving, 8 B averaging, "
                "+ packed masks); the fused kernel keeps voxel state in registers across views, so HBM traffic is "
                "far below it (see traffic) and frac may exceed 1: the kernel is FP32-issue / L1-gather bound, "
                "see DESIGN.md",
        "fused_floor_bytes_per_launch": fused_floor,
    }

    # ---- e2e: public API, pinned host masks, H2D + kernels + D2H of the int32 / float32 volume
    e2e = None
    if not args.no_e2e:
        def step_e2e():
            bp.clear()
            bp.process_views(views_host)
            return bp.get_values(gather=False) if world > 1 else bp.get_values()
        for _ in range(2):
            vol = step_e2e()
        barrier()
        moved0 = eng.ingest_bytes()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            vol = step_e2e()
        barrier()
        moved = (eng.ingest_bytes() - moved0) // args.steps  # what the library actually copied up
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": total_vv * args.steps / float(dt[0]), "unit": UNIT,
               # sharded ingest: every mask crosses PCIe once; replicated: once per rank
               "h2d_bytes_per_step": int(n_views * H * W) * (world if getattr(bp, "last_ingest", None) == "replicated" else 1),
               "ingest": getattr(bp, "last_ingest", None) or "single GPU",
               "d2h_bytes_per_step": int(nx * ny * nz * 4), "ms_per_step": 1e3 * float(dt[0]) / args.steps}
        if world == 1 and os.environ.get("BP_VIEW_RECT", "0") not in ("", "0"):
            # experimental: only the rectangle of each mask that the grid projects into crosses PCIe (bp_view_rect)
            e2e["h2d_bytes_per_step"] = int(moved)
            e2e["ingest"] = "single GPU, view rectangles"
        if world == 1 and mode == "carving" and os.environ.get("BP_SPEC_READBACK", "0") not in ("", "0"):
            # experimental: early D2H under the uploads, then only the rows that changed (bp_set_readback_target)
            patched, rows = eng.spec_stats()
            e2e["readback"] = {"kind": "speculative", "rows_patched": int(patched), "rows": int(rows)}
    else:
        step_resident()
        vol = bp.get_values(gather=False) if world > 1 else bp.get_values()

    # ---- CPU baseline (rank 0, N=1 only): oracle restatement on a bounded sample, also used as a checker
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu and with_cpu:
        n_pl = args.cpu_planes or auto_planes(w, mode)
        planes = sample_planes(nx, n_pl)
        gpu_planes = [np.array(vol[p - i0]) for p in planes]
        v, t, cores, ok = cpu_leg(w, masks_np, mode, planes, "port", gpu_planes)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{len(planes)} of {nx} axis-0 planes (evenly spread) x all {n_views} views, {t:.1f} s; "
                         f"oracle/bp_oracle.c, OpenMP over {cores} threads",
               "sample_matches_gpu": ok}

    work = None
    if mode == "carving" and hasattr(eng, "work_stats"):
        step_resident()
        eng.sync()
        eng.work_stats()  # reset the counter, then count exactly one step
        step_resident()
        ev, blocks, dead = eng.work_stats()
        work = {"nominal_voxel_views": float(len(eng.planes)) * ny * nz * n_views, "evaluated_voxel_views": ev or None,
                "cull_blocks": blocks, "cull_blocks_dead": dead}
    bp.close()
    return {"value": value, "ms_per_step": ms_step, "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
            "gpu_launches": int(tsum[3]), "clocks": clk, "work": work}


# ---------------------------------------------------------------------------------------------
# the B200 arm
# ---------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from plant3dvision_b200 import synth
    from plant3dvision_b200.cl import Backprojection

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        from plant3dvision_b200.parallel import bind_to_gpu_numa_node
        bind_to_gpu_numa_node(local)  # before any pinned allocation: host buffers land on the GPU's own socket
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    w = synth.workload(args.workload, args.scale)
    nx, ny, nz = w["shape"]
    n_views = len(w["cams"])
    H, W = w["H"], w["W"]
    i0, i1 = rank * nx // world, (rank + 1) * nx // world
    cams = w["cams"].astype(np.float32)

    # ---- inputs: rendered once, kept (a) pinned on the host for e2e, (b) resident in HBM for `value`
    masks_np = render_all(w, dev, args.mode, pinned=True)
    masks_d = torch.from_numpy(masks_np).to(dev)
    ptrs = [masks_d[v].data_ptr() for v in range(n_views)]
    hw = np.tile(np.array([[H, W]], dtype=np.int32), (n_views, 1))
    views_host = [(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], masks_np[v]) for v in range(n_views)]

    m = measure_mode(args, args.mode, w, cams, masks_np, ptrs, hw, views_host, world, rank, local, dev)
    # BASELINE.json's metric covers both kernels: the default (carving) run also measures averaging on the
    # same views (the 0/255 masks read as probability maps 0.0 / 1.0), reported under "averaging"
    other = None
    if args.mode == "carving" and not args.no_averaging:
        a = measure_mode(args, "averaging", w, cams, masks_np, ptrs, hw, views_host, world, rank, local, dev)
        other = {"metric": METRIC, "value": a["value"], "unit": UNIT, "ms_per_step": a["ms_per_step"], "dtype": "f32",
                 "e2e": a["e2e"], "roofline": a["roofline"], "cpu_baseline": a["cpu_baseline"],
                 "gpu_launches": a["gpu_launches"],
                 "workload": "same grid and views, type='averaging', log=False, uint8 maps"}
    value, ms_step, roofline, e2e, cpu, clk = (m["value"], m["ms_per_step"], m["roofline"], m["e2e"],
                                               m["cpu_baseline"], m["clocks"])
    launches_total = m["gpu_launches"] + (other["gpu_launches"] if other else 0)
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32 projection, i8 state" if args.mode == "carving" else "f32",
            "data": "synthetic", "config": workload_config(w, args, world),
            "e2e": e2e, "gpu_launches": launches_total, "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
            "work": m["work"],
        }
        if other is not None:
            line["averaging"] = other
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    # stdout carries exactly ONE JSON line: native libraries that print to file descriptor 1 (NCCL's version banner)
    # are pointed at stderr, Python's own sys.stdout keeps the original descriptor
    _out = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(_out, "w")
    a = parse_args()
    sys.exit(run_reference(a) if a.impl == "reference" else run_b200(a))

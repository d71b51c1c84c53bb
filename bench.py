#!/usr/bin/env python3
"""bench.py — voxel-views/s of the back-projection hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload C5|C1|C1t|C2|C3|C4] [--mode carving|averaging]
                    [--sharding stripes|slabs|views] [--scale S]

Headline (default, every N): C5 of BASELINE.md §3 - carving a 1024^3 grid from 240 views of 6000x4000 masks (the
configuration north_star's target is quoted on; 1 GiB of int8 state + 5.8 GB of masks fit one B200).  With N ranks the
planes of axis 0 are dealt to the ranks in stripes of 8, every rank sees all views for its planes, no
collective on the voxel data: total work is fixed ("scaling": "strong").

One step = clear the grid + ingest every mask + the fused kernel(s) over all views.
  value : whole-job voxel-views/s with the uint8 masks already resident in HBM, device-timed (CUDA events on the
          library's stream, max over ranks)
  e2e   : the same step through the public Backprojection API from HOST masks, including the H2D copies and the
          get_values() D2H of the int32 / float32 volume, wall clock between synchronisations
  roofline, cpu_baseline, clocks, gpu_launches, work : DESIGN.md §5

The default run also carries, as sub-objects of the one JSON line,
  "averaging"     the same C5 views through type='averaging' (BASELINE.json's metric names both kernels),
  "configs"       N = 1: C1 (geom_pipe_virtual.toml: the reference's virtual_plant fixture, voxel_size 0.06, carving,
                  invert) and C3 (ml_pipe_virtual.toml: its 6 label channels, averaging through process_fileset) - both
                  checked against the committed golden hashes of the reference's own kernel source,
  "view_sharded"  N > 1: C4 (512^3 x 120 views) with every rank holding a full replica and a share of the views, the
                  NCCL MIN all-reduce of the int8 state INSIDE the timed step and timed on its own.
Every leg compares sampled planes of the GPU volume with the CPU oracle (rank 0, the planes it owns), at every N.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "voxel-views/sec"
UNIT = "voxel-views/s"
CHANNELS = ["background", "flower", "fruit", "leaf", "pedicel", "stem"]
N_SM = 148
# FP32 instructions the reference's arithmetic costs per voxel-view (kernels/backprojection.c:11-31 with the partial
# sums r*x + r*y shared along a row and the two divisions sharing one reciprocal): 9 FADD, 7 FMUL, 6 FFMA, 1 MUFU,
# 2 F2I.  The issue roofline counts only these as useful work.
PRESCRIBED_INSTR = 25.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C5", choices=["C5", "C1", "C1t", "C2", "C3", "C4"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink grid and images of a synthetic workload (debug)")
    ap.add_argument("--mode", default=None, choices=["carving", "averaging"], help="default: the workload's own")
    ap.add_argument("--sharding", default=None, choices=["stripes", "slabs", "views"], help="N > 1 (default stripes)")
    ap.add_argument("--cpu-planes", type=int, default=0, help="planes in the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / oracle-check legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-averaging", action="store_true", help="skip the secondary averaging leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs / view_sharded sub-objects")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# static evidence: peaks, ncu captures (tied to the device code they measured)
# ---------------------------------------------------------------------------------------------
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def kernel_hash():
    try:
        import __graft_entry__ as g
        return g.built_hashes()[1]
    except Exception:  # noqa: BLE001
        return None


def ncu_evidence(mode):
    """Counters of the committed `ncu --set full` capture of the dominant kernel (profiles/traffic.json), or a stub
    saying it is stale: a capture is only printed when it was taken from the device code this library is built from."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            t = json.load(f)
    except (OSError, ValueError):
        return None
    entry = (t.get("ncu") or {}).get(mode)
    if not entry:
        return None
    have = kernel_hash()
    if entry.get("kern_hash") != have:
        return {"stale": True, "captured_kern_hash": entry.get("kern_hash"), "built_kern_hash": have,
                "source": entry.get("source")}
    return dict(entry, stale=False)


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region, sampled every ~5 ms through NVML in a thread (same
    counters as the nvidia-smi clocks line of B200_PROFILING.md; nvidia-smi itself cannot sample a 100 ms region)."""

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples, self.reasons, self.power = [], set(), []
        self.max_mhz = None
        self._stop = False
        self._t = None
        self._nv = None

    def _run(self):
        nv, h = self._nv, self._h
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self._stop:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def start(self):
        try:
            import pynvml as nv
            import threading
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = self.idx
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                phys = int(vis.split(",")[self.idx])
            self._h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM))
            self._nv = nv
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        except Exception:  # noqa: BLE001
            self._nv = None
        return self

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


# ---------------------------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------------------------
def load_workload(name, scale):
    """dict(name, shape, origin, voxel_size, cams [n,16] float64, H, W, mode, log, invert, labels, source, desc);
    the masks come from `workload_masks`."""
    from plant3dvision_b200 import synth
    if name in ("C5", "C2", "C4"):
        w = synth.workload(name, scale)
        w.update(mode="carving", log=False, invert=False, labels=None, source="synthetic",
                 desc="synthetic phantom silhouettes")
        return w
    z = np.load(os.path.join(ROOT, "tests", "golden", "virtual_plant_inputs.npz"))
    bbox = ((z["bbox"][0][0], z["bbox"][0][1]), (z["bbox"][1][0], z["bbox"][1][1]), (z["bbox"][2][0], z["bbox"][2][1]))
    if name in ("C1", "C1t"):  # configs/geom_pipe_virtual.toml:6 / test_geom_pipe_virtual.toml:14
        vs = 0.06 if name == "C1" else 0.1
        mode, log, invert, labels = "carving", False, True, None
    else:                      # C3: configs/ml_pipe_virtual.toml:19-26
        vs, mode, log, invert, labels = 0.3, "averaging", False, False, list(CHANNELS)
    shape, origin = synth.shape_origin(bbox, vs)
    return dict(name=name, bbox=bbox, voxel_size=vs, shape=shape, origin=origin, cams=z["cams"], H=896, W=896,
                mode=mode, log=log, invert=invert, labels=labels, source="reference fixture",
                desc="the reference's tests/testdata/virtual_plant fixture (tests/golden/virtual_plant_inputs.npz)",
                _maps={c: z[f"map_{c}"] for c in CHANNELS})


def workload_masks(w, dev, mode, pinned):
    """{label or None: uint8 [n, H, W] numpy (pinned when it feeds the e2e leg)}."""
    import torch
    from plant3dvision_b200 import synth
    if "_maps" in w:
        labels = w["labels"] or ["background"]
        out = {}
        for lab in labels:
            t = torch.from_numpy(np.ascontiguousarray(w["_maps"][lab]))
            out[lab if w["labels"] else None] = (t.pin_memory() if pinned and torch.cuda.is_available() else t).numpy()
        return out
    n, H, W = len(w["cams"]), w["H"], w["W"]
    host = torch.empty((n, H, W), dtype=torch.uint8, pin_memory=(pinned and torch.cuda.is_available()))
    soft = 2.0 if mode == "averaging" else 0.0
    for a in range(0, n, 16):
        b = min(n, a + 16)
        host[a:b].copy_(synth.render_masks(w["cams"][a:b], w["capsules"], H, W, device=dev, soft_px=soft))
    return {None: host.numpy()}


def workload_config(w, mode, n_gpus, sharding, scale):
    nl = len(w["labels"]) if w["labels"] else 1
    how = {"stripes": f"axis-0 stripes (blocks of 8 planes dealt round robin) x{n_gpus}; no collective on the voxel data",
           "slabs": f"contiguous axis-0 slabs x{n_gpus}; no collective on the voxel data",
           "views": f"view subsets x{n_gpus} on full replicas + one NCCL all-reduce (MIN int8 / SUM f32) per label"}
    big = w["name"] in ("C5", "C4", "C2", "C1", "C1t") and scale >= 0.5
    return {
        "workload": f"{w['name']} {mode}: grid {w['shape'][0]}x{w['shape'][1]}x{w['shape'][2]}, "
                    f"{len(w['cams'])} views of {w['W']}x{w['H']} uint8 masks"
                    f"{' x %d labels' % nl if nl > 1 else ''} ({w['desc']})",
        "grid": [int(s) for s in w["shape"]], "views": len(w["cams"]), "labels": nl, "mask_hw": [w["H"], w["W"]],
        "mode": mode, "invert": bool(w["invert"]), "log": bool(w["log"]),
        "sharding": how[sharding] if n_gpus > 1 else "single GPU",
        "l2": "inputs larger than L2 (masks + grid >> 126 MB); grid re-seeded every step" if big
              else "small workload: fits L2 (the reference's own configuration / a debug scale)",
        "scale": scale,
    }


class MemFile:
    """In-memory stand-in for a plantdb File (duck-typed like fsdb_lite.File): the e2e leg of the fixture workloads
    goes through Backprojection.process_fileset, whose metric excludes the image decode from disk."""

    def __init__(self, fid, channel, cam, image):
        self.id = fid
        self._md = {"channel": channel, "camera": cam}
        self._im = image

    def get_metadata(self, key=None, default=None):
        return self._md if key is None else self._md.get(key, default)

    def read_image(self):
        return self._im


def mem_fileset(w, masks):
    files = []
    for v in range(len(w["cams"])):
        c = np.asarray(w["cams"][v], dtype=np.float64)
        cam = {"camera_model": {"params": c[0:4].tolist() + [0.0] * 4}, "rotmat": c[4:13].reshape(3, 3).tolist(),
               "tvec": c[13:16].tolist()}
        for lab, m in masks.items():
            files.append(MemFile(f"{v:05d}_{lab or 'background'}", lab or "background", cam, m[v]))
    return files


# ---------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/)
# ---------------------------------------------------------------------------------------------
def sample_planes(candidates, n):
    """`n` evenly spread planes out of the (sorted) global plane indices `candidates`."""
    candidates = list(candidates)
    n = max(1, min(n, len(candidates)))
    return sorted({candidates[int((q + 0.5) * len(candidates) / n)] for q in range(n)})


def auto_planes(w, mode, long_sample=False):
    """Planes of the CPU sample.  `long_sample` (the timed cpu_baseline of the single-GPU line): about 10 s of work on
    all cores - carving skips carved voxels early (port: ~16 G vv/s on 16 threads), averaging touches every voxel-view
    (~1.5 G vv/s); otherwise (the oracle check at N > 1, where torchrun leaves one OpenMP thread per rank) a second or two."""
    per_plane = w["shape"][1] * w["shape"][2] * len(w["cams"]) * (len(w["labels"]) if w["labels"] else 1)
    if long_sample:
        budget, cap = (1.6e11, 640) if mode == "carving" else (1.5e10, 64)
    else:
        budget, cap = (2.4e10, 32) if mode == "carving" else (6e9, 32)
    return int(max(2, min(cap, budget // max(per_plane, 1))))


def _cpu_lut(log):
    from oracle import backprojection_oracle as bo
    lut = bo.img_as_float32(np.arange(256, dtype=np.uint8))
    if log:
        lut = np.log(bo.EPS + lut)
    return np.ascontiguousarray(lut, dtype=np.float32)


def port_planes(w, masks_np, mode, planes, threads=None):
    """oracle/bp_oracle.c on `planes` (each a one-plane slab with the GLOBAL plane index; OpenMP over its rows) x all
    views, uint8 masks read through the `!= 0` test / the 256-entry table - the same conversions cl.py:205-215,
    :300-301 make.  Returns (list of [ny, nz] planes, seconds, threads used)."""
    import ctypes
    from oracle import backprojection_oracle as bo
    lib = bo._lib("restatement")
    lib.oracle_max_threads.restype = ctypes.c_int
    if threads:
        lib.oracle_set_threads(int(threads))
    used = int(lib.oracle_max_threads())
    F, I, B = (ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_uint8))
    cams = np.ascontiguousarray(w["cams"], dtype=np.float32)
    ny, nz = int(w["shape"][1]), int(w["shape"][2])
    origin = np.array(w["origin"], dtype=np.float32)
    vs = ctypes.c_float(float(np.float32(w["voxel_size"])))
    shape1 = np.array([1, ny, nz], dtype=np.int32)
    lut = _cpu_lut(w["log"]) if mode == "averaging" else None
    if w["invert"]:
        masks_np = np.invert(masks_np)  # cl.py:300-301 (outside the timed region: the GPU leg's inversion is free too)
    out, t_total = [], 0.0
    for i0 in planes:
        state = np.zeros((1, ny, nz), dtype=np.int32 if mode == "carving" else np.float32)
        t0 = time.perf_counter()
        for v in range(len(cams)):
            m, c = masks_np[v], cams[v]
            H, W = m.shape
            if mode == "carving":
                lib.oracle_carve_view(m.ctypes.data_as(B), H, W, state.ctypes.data_as(I), c[0:4].ctypes.data_as(F),
                                      c[4:13].ctypes.data_as(F), c[13:16].ctypes.data_as(F), origin.ctypes.data_as(F), vs,
                                      shape1.ctypes.data_as(I), int(i0))
            else:
                lib.oracle_average_view_lut(m.ctypes.data_as(B), lut.ctypes.data_as(F), H, W, state.ctypes.data_as(F),
                                            c[0:4].ctypes.data_as(F), c[4:13].ctypes.data_as(F),
                                            c[13:16].ctypes.data_as(F), origin.ctypes.data_as(F), vs,
                                            shape1.ctypes.data_as(I), int(i0))
        t_total += time.perf_counter() - t0
        out.append(state[0])
    return out, t_total, used


def oracle_check(w, masks, mode, planes, gpu_planes_of, rtol=0.0):
    """Compare sampled planes of the GPU volume(s) with the oracle port, label by label.  gpu_planes_of(label_index)
    -> list of [ny, nz] arrays for `planes`.  Returns (all equal, seconds, threads, voxel-views computed)."""
    ok, t_total, used, vv = True, 0.0, 1, 0.0
    for li, (lab, m) in enumerate(masks.items()):
        want, t, used = port_planes(w, m, mode, planes)
        got = gpu_planes_of(li)
        for a, b in zip(got, want):
            if rtol:
                ok = ok and bool(np.allclose(a, b, rtol=rtol, atol=1e-6))
            else:
                ok = ok and bool(np.array_equal(a, b))
        t_total += t
        vv += float(len(planes)) * w["shape"][1] * w["shape"][2] * len(w["cams"])
    return ok, t_total, used, vv


def thread_scaling(w, masks_np, mode, planes):
    """Measured 1-thread -> all-thread speed-up of the port on a small sample (the figure VERDICT r1 asked for)."""
    cores = os.cpu_count() or 1
    _, t1, _ = port_planes(w, masks_np, mode, planes[:1], threads=1)
    _, tn, used = port_planes(w, masks_np, mode, planes[:1], threads=cores)
    return {"threads": used, "t_1_thread_s": t1, "t_all_threads_s": tn, "speedup": t1 / tn if tn > 0 else None}


def reference_sample(w, masks_np, mode, planes):
    """The reference's own kernel source (oracle/_ref, compiled unmodified from kernels/backprojection.c) the way
    process_view drives it (cl.py:190-227): per view ONE widening of the mask to int32 / float32 (cl.py:205-215), then
    the kernel - here over the sampled planes only (one-plane grids whose origin is shifted to the plane; the float32
    world point is the same).  Returns (kernel seconds, widening seconds over all views, threads)."""
    import ctypes
    from oracle import backprojection_oracle as bo
    lib = bo._lib("ref_off")
    F, I = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32)
    cams = np.ascontiguousarray(w["cams"], dtype=np.float32)
    ny, nz = int(w["shape"][1]), int(w["shape"][2])
    origin = np.array(w["origin"], dtype=np.float32)
    vs = np.float32(w["voxel_size"])
    shape1 = np.array([1, ny, nz], dtype=np.int32)
    states = [np.zeros((1, ny, nz), dtype=np.int32 if mode == "carving" else np.float32) for _ in planes]
    volinfos = [np.array([origin[0] + np.float32(i0) * vs, origin[1], origin[2], vs], dtype=np.float32) for i0 in planes]
    t_widen = t_kernel = 0.0
    for v in range(len(cams)):
        m, c = masks_np[v], cams[v]
        H, W = m.shape
        t0 = time.perf_counter()
        if w["invert"]:
            m = np.invert(m)                                                # cl.py:300-301
        if mode == "carving":
            mh = np.ascontiguousarray(m, dtype=np.int32)                    # cl.py:215
        else:
            mf = bo.img_as_float32(m)                                       # cl.py:205-206
            if w["log"]:
                mf = np.log(bo.EPS + mf)                                    # cl.py:207-208
            mh = np.ascontiguousarray(mf, dtype=np.float32)
        t1 = time.perf_counter()
        for st, vi in zip(states, volinfos):
            if mode == "carving":
                lib.ref_carve_view(mh.ctypes.data_as(I), H, W, st.ctypes.data_as(I), c[0:4].ctypes.data_as(F),
                                   c[4:13].ctypes.data_as(F), c[13:16].ctypes.data_as(F), vi.ctypes.data_as(F),
                                   shape1.ctypes.data_as(I))
            else:
                lib.ref_average_view(mh.ctypes.data_as(F), H, W, st.ctypes.data_as(F), c[0:4].ctypes.data_as(F),
                                     c[4:13].ctypes.data_as(F), c[13:16].ctypes.data_as(F), vi.ctypes.data_as(F),
                                     shape1.ctypes.data_as(I))
        t2 = time.perf_counter()
        t_widen += t1 - t0
        t_kernel += t2 - t1
    return t_kernel, t_widen, int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1))


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only).  Each step is
    a bounded sample - P planes x all views - and the per-view mask widening, which the reference pays ONCE per view
    for all nx planes, is charged to the sample in proportion (P / nx of it), so that value = what the full job would
    sustain.  Both raw times are printed."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    import torch
    from oracle import backprojection_oracle as bo
    w = load_workload(args.workload, args.scale)
    mode = args.mode or w["mode"]
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    masks = workload_masks(w, dev, mode, pinned=False)
    m0 = next(iter(masks.values()))
    kind = "reference" if bo.have_backend("ref_off") else "port"
    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    nx = int(w["shape"][0])
    n_pl = args.cpu_planes or max(2, auto_planes(w, mode) // 4)
    planes = sample_planes(range(nx), n_pl)
    per_sample_vv = float(len(planes)) * w["shape"][1] * w["shape"][2] * len(w["cams"])

    def one_step():
        if kind == "reference":
            tk, tw, thr = reference_sample(w, m0, mode, planes)
            return tk + tw * len(planes) / nx, tk, tw, thr
        _, t, thr = port_planes(w, m0, mode, planes)
        return t, t, 0.0, thr

    for _ in range(min(args.warmup, 1)):
        one_step()
    eff, tks, tws, thr = [], [], [], cores
    for _ in range(args.steps):
        e, tk, tw, thr = one_step()
        eff.append(e)
        tks.append(tk)
        tws.append(tw)
    value = per_sample_vv * len(eff) / sum(eff)
    sample = (f"{len(planes)} of {nx} axis-0 planes (evenly spread) x all {len(w['cams'])} views per step; "
              f"{'oracle/_ref = kernels/backprojection.c compiled unmodified' if kind == 'reference' else 'oracle/bp_oracle.c'}, "
              f"OpenMP over {thr} threads; kernel {np.mean(tks):.2f} s + widening of all masks {np.mean(tws):.2f} s per step, "
              f"the widening charged {len(planes)}/{nx} (the reference widens once per view for all planes, cl.py:215)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(eff) / len(eff),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 projection, i32 state" if mode == "carving" else "f32", "data": w["source"],
        "config": workload_config(w, mode, 1, "stripes", args.scale),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": thr, "kind": kind, "sample": sample,
                         "kernel_s_per_step": float(np.mean(tks)), "widen_all_masks_s_per_step": float(np.mean(tws))},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
# the B200 arm
# ---------------------------------------------------------------------------------------------
class Ctx:
    pass


def measure(args, ctx, w, mode, masks, sharding, with_cpu=True, with_e2e=True, steps=None):
    """One workload x mode on the already rendered masks: device-resident throughput, roofline, e2e, oracle check."""
    import torch
    import torch.distributed as dist
    from plant3dvision_b200.cl import Backprojection
    world, rank, local, dev = ctx.world, ctx.rank, ctx.local, ctx.dev
    steps = steps or args.steps
    nx, ny, nz = [int(s) for s in w["shape"]]
    cams = np.ascontiguousarray(w["cams"], dtype=np.float32)
    n_views, H, W = len(cams), w["H"], w["W"]
    labels = list(masks.keys())
    n_labels = len(labels)
    if world > 1:
        from plant3dvision_b200.parallel import DistributedBackprojection
        kw = {"stripe_lg": int(os.environ.get("BP_STRIPE_LG", "3"))} if sharding == "stripes" else {}
        bp = DistributedBackprojection(w["shape"], w["origin"], w["voxel_size"], type=mode, log=w["log"],
                                       labels=w["labels"], sharding=sharding, device=local, **kw)
        eng = bp.local
    else:
        bp = eng = Backprojection(w["shape"], w["origin"], w["voxel_size"], type=mode, log=w["log"], labels=w["labels"],
                                  device=local)
    my_planes = [int(p) for p in eng.planes]
    ext = torch.cuda.ExternalStream(eng.stream(), device=dev)
    masks_d = {lab: torch.from_numpy(m).to(dev) for lab, m in masks.items()}
    ptrs = {lab: [t[v].data_ptr() for v in range(n_views)] for lab, t in masks_d.items()}
    hw = np.tile(np.array([[H, W]], dtype=np.int32), (n_views, 1))
    reduce_ms = []

    def step_resident(timed_reduce=False):
        for lab in labels:
            bp.clear()
            bp.process_device_views(cams, ptrs[lab], hw, invert=w["invert"])
            if world > 1 and sharding == "views":
                # the exchange step of view sharding belongs to the step: all-reduce of the raw device state
                if timed_reduce:
                    eng.sync()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    bp.reduce()
                    b.record()
                    torch.cuda.synchronize()
                    reduce_ms.append(a.elapsed_time(b))
                else:
                    bp.reduce()

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_resident()
    barrier()
    eng.timing()
    _, _, launches0 = eng.timing()
    clocks = ClockSampler(local).start() if rank == 0 else None
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(steps):
        step_resident()
    e1.record(ext)
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    # view sharding runs the collective on torch's stream with host synchronisation in between: wall clock there
    ms_total = wall_ms if (world > 1 and sharding == "views") else e0.elapsed_time(e1)
    clk = clocks.stop() if clocks else None
    kernel_ms, ingest_ms, launches1 = eng.timing()
    t = torch.tensor([ms_total, kernel_ms, ingest_ms, float(launches1 - launches0)], dtype=torch.float64, device=dev)
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    ms_step = float(tmax[0]) / steps
    total_vv = float(nx) * ny * nz * n_views * n_labels
    value = total_vv / (ms_step * 1e-3)
    if world > 1 and sharding == "views":
        for _ in range(2):
            step_resident(timed_reduce=True)

    # ---- work actually done by the carve kernels (counting variant, one untimed step) and the cull's statistics
    work = None
    if mode == "carving":
        os.environ["BP_COUNT_WORK"] = "1"
        try:
            cnt = Backprojection(w["shape"], w["origin"], w["voxel_size"], type=mode, device=local,
                                 **({"stripes": eng.stripes} if getattr(eng, "stripes", None) else
                                    {"i_range": eng.i_range}))
        finally:
            os.environ.pop("BP_COUNT_WORK", None)
        lab0 = labels[0]
        sel = bp.my_views(n_views) if (world > 1 and sharding == "views") else list(range(n_views))
        cnt.process_device_views(cams[sel], [ptrs[lab0][v] for v in sel], hw[sel], invert=w["invert"])
        cnt.sync()
        ev, blocks, dead = cnt.work_stats()
        cnt.close()
        nominal = float(len(my_planes)) * ny * nz * len(sel)
        work = {"nominal_voxel_views": nominal, "evaluated_voxel_views": float(ev),
                "evaluated_fraction": float(ev) / nominal if nominal else None,
                "cull_blocks": int(blocks), "cull_blocks_dead": int(dead), "scope": "rank 0, one step"}

    elif hasattr(eng, "work_stats"):
        # averaging has no early exit; view skipping replaces whole (block, view) pairs by one add
        eng.sync()
        eng.work_stats()
        step_resident()
        eng.sync()
        ev, blocks, _ = eng.work_stats()
        nominal = float(len(my_planes)) * ny * nz * n_views * n_labels
        work = {"nominal_voxel_views": nominal, "evaluated_voxel_views": float(ev),
                "evaluated_fraction": float(ev) / nominal if nominal else None, "skip_blocks": int(blocks),
                "scope": "this rank, one step, all labels"}

    # ---- roofline of the dominant kernel(s): the binding resource is the FP32 issue rate, not HBM
    peak_hbm, peak_src = peaks()
    k_ms = float(tmax[1]) / steps / n_labels      # per launch group (all views of one label)
    n_local = len(my_planes)
    evaluated = float(n_local) * ny * nz * n_views
    if work:
        evaluated = work["evaluated_voxel_views"] / (n_labels if mode == "averaging" else 1)
    sm_hz = (clk["sm_mhz"] if clk and clk.get("sm_mhz") else 1965.0) * 1e6
    issue_peak = N_SM * 4 * sm_hz / 1e9           # G warp-instructions / s
    useful = evaluated * PRESCRIBED_INSTR / 32.0  # warp-instructions of prescribed arithmetic, this rank, one label
    achieved = useful / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    bytes_per_vv = 4.0 if mode == "carving" else 8.0
    mask_bytes = n_views * (H * W / 8.0 if mode == "carving" else float(H * W))
    streaming = bytes_per_vv * float(n_local) * ny * nz * n_views + mask_bytes
    fused_floor = (1.0 if mode == "carving" else 8.0) * float(n_local) * ny * nz + mask_bytes
    ncu = ncu_evidence(mode) if w["name"] == "C5" and world == 1 else None
    traffic = ncu.get("dram_bytes") if ncu and not ncu.get("stale") else None
    gbps = (lambda b: b / (k_ms * 1e-3) / 1e9 / peak_hbm) if k_ms > 0 else (lambda b: None)
    roofline = {
        "bound": "issue",
        "kernel": "carve_items_kernel + carve_segments_kernel (+ cull_blocks / classify_items / decode_items / tile_any / tile_sat)" if mode == "carving"
                  else "average_kernel<u8> (+ map_any / tile_sat / avg_skip)",
        "achieved": achieved, "peak": issue_peak, "unit": "G warp-instr/s", "frac": achieved / issue_peak,
        "definition": f"achieved = voxel-view projections the kernels evaluated "
                      f"({'counted by the BP_COUNT_WORK variant' if mode == 'carving' else 'all voxel-views minus the ones view skipping replaced by one add'}) "
                      f"x {PRESCRIBED_INSTR:.0f} FP32 instructions the reference's arithmetic costs per voxel-view / 32 lanes / "
                      f"kernel time; peak = {N_SM} SMs x 4 schedulers x SM clock under load",
        "traffic": traffic, "kernel_ms_per_launch": k_ms, "ingest_ms_per_step": float(tmax[2]) / steps,
        "hbm": {
            "peak_GBps": peak_hbm, "peak_source": peak_src,
            "streaming_equivalent_bytes": streaming, "streaming_equivalent_frac": gbps(streaming),
            "fused_floor_bytes": fused_floor, "fused_floor_frac": gbps(fused_floor),
            "hbm_frac_physical": gbps(traffic) if traffic else None,
            "note": "streaming_equivalent = SURVEY 8(d) algorithmic bytes of the reference's per-view formulation (4 B / "
                    "voxel-view carving, 8 B averaging, + masks): > 1 by construction for a fused kernel that keeps the "
                    "voxel state in registers and skips carved voxels - NOT a roofline fraction; fused_floor = one pass "
                    "over the state + every mask once; hbm_frac_physical = ncu DRAM bytes / kernel time / peak",
        },
        "ncu": ncu,
    }

    # ---- e2e: public API from HOST masks, H2D + kernels (+ collective) + D2H of the volume
    e2e, vols = None, None
    if with_e2e:
        if w["labels"] or "_maps" in w:
            files = mem_fileset(w, masks)

            def step_e2e():
                bp.clear()  # (process_fileset clears BETWEEN labels, cl.py:249-252: a fresh grid is the caller's job)
                if world > 1:
                    return bp.process_fileset(files, "camera", invert=w["invert"], gather=False)
                return bp.process_fileset(files, "camera", invert=w["invert"])
            api = "clear(); Backprojection.process_fileset(files, 'camera', invert=%s)" % w["invert"]
        else:
            m0 = masks[None]
            views_host = [(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], m0[v]) for v in range(n_views)]

            def step_e2e():
                bp.clear()
                bp.process_views(views_host)
                return bp.get_values(gather=False) if world > 1 else bp.get_values()
            api = "clear(); process_views(views); get_values()"
        for _ in range(2):
            vols = step_e2e()
        barrier()
        moved0 = eng.ingest_bytes() + getattr(bp, "h2d_torch_bytes", 0)
        t0 = time.perf_counter()
        for _ in range(steps):
            vols = step_e2e()
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        moved = torch.tensor([float(eng.ingest_bytes() + getattr(bp, "h2d_torch_bytes", 0) - moved0) / steps], dtype=torch.float64, device=dev)
        # int32 / float32 volumes; a label stack leaves the device as float64 (the tail kernel widens, cl.py:249)
        d2h = torch.tensor([float(np.asarray(vols).size) * (8 if np.asarray(vols).dtype == np.float64 else 4)],
                           dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            dist.all_reduce(moved, op=dist.ReduceOp.SUM)
            dist.all_reduce(d2h, op=dist.ReduceOp.SUM)
        e2e = {"value": total_vv * steps / float(dt[0]), "unit": UNIT, "ms_per_step": 1e3 * float(dt[0]) / steps,
               "h2d_bytes_per_step": int(moved[0]), "d2h_bytes_per_step": int(d2h[0]), "api": api,
               "h2d_note": "bytes copied up, all ranks (bp_ingest_bytes + the sharded averaging ingest's own torch copies): view "
                           "rectangles trim each mask to the pixels the grid projects into",
               "ingest": getattr(bp, "last_ingest", None) or "single GPU"}
        if world == 1 and mode == "carving" and getattr(eng, "speculative_readback", False):
            patched, rows = eng.spec_stats()
            e2e["readback"] = {"kind": "speculative", "rows_patched": int(patched), "rows": int(rows)}
    else:
        step_resident()
        if n_labels == 1:
            vols = bp.get_values(gather=False) if world > 1 else bp.get_values()

    # ---- oracle check on sampled planes (rank 0, the planes it owns) + the timed CPU baseline
    cpu, check = None, None
    if rank == 0 and with_cpu and not args.no_cpu and vols is not None:
        vols_np = np.asarray(vols)
        if vols_np.ndim == 3:
            vols_np = vols_np[None]
        n_pl = args.cpu_planes or auto_planes(w, mode, long_sample=(world == 1))
        planes = sample_planes(my_planes, n_pl)
        idx = [my_planes.index(p) for p in planes]
        if world > 1 and sharding == "views":
            idx = planes  # full replica
        rtol = 1e-5 if (mode == "averaging" and world > 1 and sharding == "views") else 0.0
        if mode == "averaging" and vols_np.dtype == np.float64:
            def gp(li):
                return [vols_np[li, q].astype(np.float32) for q in idx]   # exact: the float64 stack holds float32 sums
        else:
            def gp(li):
                return [vols_np[li, q] for q in idx]
        ok, t_cpu, thr, vv = oracle_check(w, masks, mode, planes, gp, rtol=rtol)
        check = {"sample_matches_gpu": ok, "planes": len(planes), "labels": n_labels,
                 "bar": "bit-exact" if not rtol else "1e-5 relative (summation order differs)"}
        if world == 1:
            m0 = next(iter(masks.values()))
            cpu = {"value": vv / t_cpu, "unit": UNIT, "cores": thr, "kind": "port",
                   "sample": f"{len(planes)} of {nx} axis-0 planes (evenly spread) x all {n_views} views"
                             f"{' x %d labels' % n_labels if n_labels > 1 else ''}, {t_cpu:.1f} s; oracle/bp_oracle.c, "
                             f"OpenMP over (plane, row) pairs, {thr} threads",
                   "thread_scaling": thread_scaling(w, m0, mode, planes), "sample_matches_gpu": ok}

    bp.close()
    out = {"value": value, "ms_per_step": ms_step, "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
           "oracle_check": check, "gpu_launches": int(tsum[3]), "clocks": clk, "work": work}
    if reduce_ms:
        r = torch.tensor([float(np.mean(reduce_ms))], dtype=torch.float64, device=dev)
        dist.all_reduce(r, op=dist.ReduceOp.MAX)
        state_bytes = float(nx) * ny * ((nz + 15) // 16 * 16) * (1 if mode == "carving" else 4)
        out["allreduce"] = {"ms": float(r[0]), "bytes": state_bytes, "op": "MIN int8" if mode == "carving" else "SUM f32",
                            "algbw_GBps": state_bytes / (float(r[0]) * 1e-3) / 1e9, "inside_timed_step": True}
    return out


def golden_meta():
    try:
        z = np.load(os.path.join(ROOT, "tests", "golden", "virtual_plant_golden.npz"))
        return json.loads(str(z["meta"]))
    except Exception:  # noqa: BLE001
        return {}


def golden_check(name, local):
    """C1 / C3 volumes in full against the committed sha256 of the reference's own kernel source (tests/golden)."""
    import hashlib
    from plant3dvision_b200.cl import Backprojection
    meta = golden_meta()
    w = load_workload(name, 1.0)
    cams = np.ascontiguousarray(w["cams"], dtype=np.float32)
    res = {}
    if name in ("C1", "C1t"):
        key = f"carve_vs{w['voxel_size']}"
        if key not in meta:
            return {"golden": "absent"}
        bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="carving", device=local)
        m = w["_maps"]["background"]
        bp.process_views([(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], m[v]) for v in range(len(cams))], invert=True)
        vol = bp.get_values()
        res[key] = hashlib.sha256(np.ascontiguousarray(vol).tobytes()).hexdigest() == meta[key]["sha256"]
        bp.close()
    else:
        bp = Backprojection(w["shape"], w["origin"], w["voxel_size"], type="averaging", log=False, device=local)
        for n, lab in enumerate(CHANNELS):
            key = f"avg_vs0.3_log0_{lab}"
            if key not in meta:
                continue
            if n:
                bp.clear()
            m = w["_maps"][lab]
            bp.process_views([(cams[v, 0:4], cams[v, 4:13], cams[v, 13:16], m[v]) for v in range(len(cams))])
            res[key] = hashlib.sha256(np.ascontiguousarray(bp.get_values()).tobytes()).hexdigest() == meta[key]["sha256"]
        bp.close()
    return {"golden_sha256_match": res, "all": bool(res) and all(res.values()),
            "golden": "tests/golden/virtual_plant_golden.npz: outputs of the reference's own kernel source (oracle/_ref)"}


def sub_object(m, w, mode, ctx, sharding, scale):
    keep = {k: m[k] for k in ("value", "ms_per_step", "e2e", "roofline", "cpu_baseline", "oracle_check", "gpu_launches",
                              "work", "allreduce") if m.get(k) is not None}
    keep.update(metric=METRIC, unit=UNIT, dtype="f32 projection, i8 state" if mode == "carving" else "f32",
                config=workload_config(w, mode, ctx.world, sharding, scale), data=w["source"])
    return keep


def run_b200(args):
    import torch
    import torch.distributed as dist
    ctx = Ctx()
    ctx.world = int(os.environ.get("WORLD_SIZE", "1"))
    ctx.rank = int(os.environ.get("RANK", "0"))
    ctx.local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the B200 arm has no CPU fallback")
    torch.cuda.set_device(ctx.local)
    if ctx.world > 1:
        from plant3dvision_b200.parallel import bind_to_gpu_numa_node
        bind_to_gpu_numa_node(ctx.local)  # before any pinned allocation: host buffers land on the GPU's own socket
        dist.init_process_group("nccl", device_id=torch.device("cuda", ctx.local))
    ctx.dev = torch.device("cuda", ctx.local)
    sharding = args.sharding or "stripes"

    w = load_workload(args.workload, args.scale)
    mode = args.mode or w["mode"]
    masks = workload_masks(w, ctx.dev, mode, pinned=True)
    m = measure(args, ctx, w, mode, masks, sharding, with_e2e=not args.no_e2e)
    line = None
    if ctx.rank == 0:
        line = {
            "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": ctx.world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": m["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 projection, i8 state" if mode == "carving" else "f32", "data": w["source"],
            "config": workload_config(w, mode, ctx.world, sharding, args.scale),
            "e2e": m["e2e"], "gpu_launches": m["gpu_launches"], "clocks": m["clocks"], "roofline": m["roofline"],
            "cpu_baseline": m["cpu_baseline"], "oracle_check": m["oracle_check"], "work": m["work"],
            "library": {"kern_hash": kernel_hash()},
        }
        if "allreduce" in m:
            line["allreduce"] = m["allreduce"]
    launches = m["gpu_launches"]
    default_run = args.workload == "C5" and mode == "carving"
    # BASELINE.json's metric names both kernels: the default run also measures averaging on the same views (the 0/255
    # masks read as probability maps 0.0 / 1.0)
    if default_run and not args.no_averaging:
        a = measure(args, ctx, w, "averaging", masks, sharding, with_e2e=not args.no_e2e)
        launches += a["gpu_launches"]
        if line is not None:
            line["averaging"] = sub_object(a, w, "averaging", ctx, sharding, args.scale)
            line["averaging"]["workload"] = "same grid and views, type='averaging', log=False, uint8 maps"
    del masks
    torch.cuda.empty_cache()
    if default_run and not args.no_extra:
        if ctx.world == 1:
            # the reference's own configurations on its own fixture data (BASELINE.json configs[0] and [2]: C1, C3, checked
            # in full against the committed hashes of the reference's kernel source), and the two synthetic ones at the
            # reference's real-scan and cylinder geometry (configs[1] and [3]: C2, C4) on one GPU
            configs = {}
            for name in ("C1", "C3", "C2", "C4"):
                wx = load_workload(name, 1.0)
                mx = workload_masks(wx, ctx.dev, wx["mode"], pinned=name in ("C2", "C4"))  # decoded images are pageable arrays
                x = measure(args, ctx, wx, wx["mode"], mx, sharding, steps=max(3, args.steps))
                launches += x["gpu_launches"]
                configs[name] = sub_object(x, wx, wx["mode"], ctx, sharding, 1.0)
                if name in ("C1", "C3"):
                    configs[name]["golden"] = golden_check(name, ctx.local)
                del mx
            line["configs"] = configs
        else:
            # north_star's first partitioning: view subsets + NCCL all-reduce of the voxel grid (BASELINE.json configs[3])
            w4 = load_workload("C4", args.scale)
            m4 = workload_masks(w4, ctx.dev, "carving", pinned=True)
            x = measure(args, ctx, w4, "carving", m4, "views")
            launches += x["gpu_launches"]
            if line is not None:
                line["view_sharded"] = sub_object(x, w4, "carving", ctx, "views", args.scale)
    if line is not None:
        line["gpu_launches"] = launches
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
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

"""Multi-GPU back-projection: one process per GPU (``torchrun``), ``torch.distributed`` for the plumbing.

The reference has no multi-device path (one OpenCL context, one queue).  Two ways to
split the work of ``Backprojection.process_fileset`` (cl.py:234-305), both exact:

``sharding="slabs"``  rank r holds planes ``[r*nx/G, (r+1)*nx/G)`` of axis 0 (contiguous in the C-order
    layout, kernels/common.h:1-12) and processes EVERY view for them.  Voxels are independent, so
    there is no collective on the data path; world points use the global plane index, so every slab
    is bit-identical to the same planes of a single-GPU run.  ``get_values(gather=True)`` all-gathers.
``sharding="views"``  every rank holds a full replica and processes views ``v = rank (mod G)``; the
    replicas are then combined with ONE all-reduce of the raw device state:
      carving    element-wise MIN over the int8 state 0 carved < 1 seen < 2 unseen.  The final
                 state of a voxel is order-independent (carved if any view carves it, else seen if
                 any view sees it, kernels/backprojection.c:65-83), which is exactly that MIN.
                 (MIN over the reference's own {-1, 0, 1} coding would be wrong: min(0, 1) = 0.)
      averaging  element-wise SUM of the float32 partial sums (ranks > 0 are seeded with 0 instead
                 of ``default_value``).  The order of the float additions differs from the
                 sequential view loop: equal within 1e-5 relative, not bit-identical.

The engine that does the per-rank work is ``plant3dvision_b200.cl.Backprojection`` (CUDA); the
class takes it as a parameter only so that the sharding logic can be exercised on CPU ranks (gloo)
in the tests with a stand-in engine.
"""
import logging
import os

import numpy as np

logger = logging.getLogger(__name__)

CARVED, SEEN, UNSEEN = 0, 1, 2  # int8 device coding of the carving state (include/bp_b200.h)


def slab_range(nx, rank, world):
    """Planes ``[i0, i1)`` of axis 0 owned by ``rank``: contiguous, disjoint, covering ``[0, nx)``."""
    return (rank * nx) // world, ((rank + 1) * nx) // world


def partition_views(n_views, rank, world):
    """Indices of the views ``rank`` processes under view sharding (round robin)."""
    return list(range(rank, n_views, world))


def encode_carve_state(vol, default_value=0):
    """Reference coding (int32: -1 carved, 1 seen, ``default_value`` unseen; cl.py:173,
    kernels/backprojection.c:79-83) -> monotone int8 coding.  Host-side twin of the device layout."""
    vol = np.asarray(vol)
    out = np.full(vol.shape, UNSEEN, dtype=np.int8)
    out[vol == -1] = CARVED
    if int(default_value) == 0:
        out[vol == 1] = SEEN
    return out


def decode_carve_state(state, default_value=0):
    """Inverse of ``encode_carve_state`` (what ``bp_get_values`` does on the device)."""
    d = int(default_value)
    state = np.asarray(state)
    out = np.full(state.shape, d, dtype=np.int32)
    if d == -1:
        return out  # a grid seeded with -1 is inert (kernels/backprojection.c:65-67)
    out[state == CARVED] = -1
    out[state == SEEN] = 1 if d == 0 else d
    return out


def spread_order(cams):
    """Farthest-point order of the views' optical axes (rows 10..12 of the 16-float camera rows; a view and its antipode
    count as one direction): the order the library itself visits the views of a carving launch in (bp_process.cu).
    Starts at view 0; every next view is the one whose axis is farthest from all chosen so far."""
    cams = np.asarray(cams, dtype=np.float64).reshape(-1, 16)
    n = len(cams)
    if n == 0:
        return []
    ax = cams[:, 10:13]
    norm = np.linalg.norm(ax, axis=1, keepdims=True)
    ax = np.divide(ax, norm, out=np.zeros_like(ax), where=norm > 0)
    far = np.full(n, 2.0)
    used = np.zeros(n, dtype=bool)
    order, cur = [], 0
    for _ in range(n):
        order.append(cur)
        used[cur] = True
        far = np.minimum(far, 1.0 - np.abs(ax @ ax[cur]))
        cand = np.where(used, -1.0, far)
        cur = int(np.argmax(cand))
    return order


def carve_chunk_bounds(n, G, chunk_views):
    """View ranges of the sharded carving ingest: [0, head) with head = 32 rounded up to a multiple of the world size
    ``G``, then chunks of ``chunk_views`` (rounded down to a multiple of ``G``, at least ``G``); a last chunk shorter than
    half a chunk is merged into the one before.  Returns the increasing list of bounds, first 0, last ``n``."""
    head = min(n, -(-32 // G) * G)
    if head >= n:
        return [0, n]
    step = max(G, (int(chunk_views) // G) * G)
    bounds = [0] + list(range(head, n, step)) + [n]
    if len(bounds) > 3 and bounds[-1] - bounds[-2] < step // 2:
        del bounds[-2]
    return bounds


def bind_to_gpu_numa_node(device):
    """Pin this process (and therefore its first-touch / pinned host allocations) to the CPUs NVML reports as
    local to CUDA device ``device``.  With one rank per GPU, host<->device copies then stay on the GPU's own
    socket and PCIe root complex instead of crossing the inter-socket link.  Returns the CPU list, or None when
    NVML / affinity information is unavailable (nothing is changed then)."""
    import os
    try:
        import pynvml as nv
        nv.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        phys = int(device)
        if vis and all(x.strip().isdigit() for x in vis.split(",")):
            phys = int(vis.split(",")[int(device)])
        h = nv.nvmlDeviceGetHandleByIndex(phys)
        n_cpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:  # no NVML, no permission, not Linux: leave the scheduler alone
        return None


def _dist():
    import torch.distributed as dist
    return dist


class DistributedBackprojection(object):
    """``Backprojection`` over all ranks of a ``torch.distributed`` process group.

    Same constructor and methods as ``plant3dvision_b200.cl.Backprojection``; every rank must make
    the same calls with the same arguments (SPMD).  Extra keyword-only parameters:

    sharding : {"slabs", "views"}
    group : process group (default: the world group)
    engine : class doing one rank's work (default ``plant3dvision_b200.cl.Backprojection``)
    """

    def __init__(self, shape, origin, voxel_size, type="carving", default_value=0, labels=None, log=False,
                 *, sharding="slabs", group=None, engine=None, ingest="auto", **engine_kwargs):
        if sharding not in ("slabs", "stripes", "views"):
            raise ValueError(f"Unknown sharding {sharding!r}, valid values are 'slabs', 'stripes' or 'views'!")
        if type not in ("carving", "averaging"):
            raise ValueError(f"Unknown kernel type {type}, valid values are 'averaging' or 'carving'!")
        dist = _dist()
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.sharding = sharding
        self.shape = [int(s) for s in shape]
        self.origin = origin
        self.voxel_size = voxel_size
        self.type = type
        self.default_value = default_value
        self.labels = labels
        self.log = log
        self.dtype = np.int32 if type == "carving" else np.float32
        if engine is None:
            from .cl import Backprojection as engine
        local_default = default_value
        if sharding == "slabs":
            self.i_range = slab_range(self.shape[0], self.rank, self.world)
            engine_kwargs["i_range"] = self.i_range
            self.planes = np.arange(self.i_range[0], self.i_range[1])
        elif sharding == "stripes":
            # blocks of 2^stripe_lg planes dealt round robin: a plant sits in the middle of the grid, so with
            # contiguous slabs the carve kernel of the middle ranks runs 2x longer than the mean (measured at
            # G = 8, profiles/r01_bench_n8_v9.json); same exactness argument as slabs
            lg = int(engine_kwargs.pop("stripe_lg", 3))
            self._stripe_lg = lg
            engine_kwargs["stripes"] = (lg, self.world, self.rank)
            self.planes = np.array([i for i in range(self.shape[0]) if (i >> lg) % self.world == self.rank])
            self.i_range = (int(self.planes[0]) if len(self.planes) else 0,
                            int(self.planes[-1]) + 1 if len(self.planes) else 0)
        else:
            self.i_range = (0, self.shape[0])
            self.planes = np.arange(self.shape[0])
            if type == "averaging" and self.rank != 0:
                local_default = 0  # the seed must enter the all-reduce SUM exactly once
        self.local = engine(self.shape, origin, voxel_size, type=type, default_value=local_default, labels=labels,
                            log=log, **engine_kwargs)
        self._reduced = True  # nothing accumulated yet
        # view-sharded averaging: after a SUM all-reduce EVERY rank holds the total; before more views are added the
        # ranks != 0 must start again from 0, or the next reduce would count the total world-size times
        self._holds_total = False
        if ingest not in ("auto", "sharded", "replicated"):
            raise ValueError(f"Unknown ingest {ingest!r}")
        # slab sharding needs every mask on every rank.  "sharded": a rank uploads (PCIe) and bit-packs only
        # the views v = rank (mod G) and the ranks exchange the results with ONE all-gather over NVLink
        # (carving: 1 bit per pixel; averaging: the uint8 maps) instead of G copies of every mask crossing
        # PCIe.  "replicated": every rank ingests every mask itself (what happens on CPU test engines).
        self.ingest = ingest
        self._xbuf = None
        self.last_ingest = None  # how the last process_views call moved the masks
        self.trace = None  # see _tick
        self.h2d_torch_bytes = 0  # host -> device bytes this object copied through torch (the library counts its own)

    # -- work distribution ---------------------------------------------------------------------
    def my_views(self, n_views):
        if self.sharding == "views":
            return partition_views(n_views, self.rank, self.world)
        return list(range(n_views))

    # -- sharded ingest (slabs) ----------------------------------------------------------------
    def _can_shard_ingest(self, shapes, dtypes):
        if self.sharding == "views" or self.world == 1 or self.ingest == "replicated":
            return False
        ok = (hasattr(self.local, "pack_views") and len(set(shapes)) == 1 and
              all(np.dtype(d) == np.uint8 for d in dtypes))
        if ok and self.ingest == "auto" and self.type == "averaging" and self.world < 4:
            # measured at G = 2 (profiles/r01_bench_n2_*.json): the averaging kernel hides a full upload per rank
            # (225 ms replicated vs 250 ms sharded); the exchange pays off once PCIe time exceeds kernel time / G
            return False
        if not ok and self.ingest == "sharded":
            raise ValueError("sharded ingest needs a CUDA engine and uint8 masks of one size")
        return ok

    #: raw mask bytes per pipeline chunk of the sharded ingest
    CHUNK_BYTES = 512 << 20
    #: views per chunk of the sharded carving ingest after the first 32
    CARVE_CHUNK_VIEWS = 48

    def _tick(self, label):
        """Diagnostics (``self.trace = []`` switches it on; never during a timed run): synchronise the device and note the
        wall clock, so that the phases of the sharded ingest can be timed one by one, un-overlapped (tools/e2e_trace.py)."""
        if self.trace is not None:
            import time
            import torch
            torch.cuda.synchronize()
            self.trace.append((label, time.perf_counter()))

    def _exchange_buffers(self, n_per, stride):
        import torch
        want = (self.world, n_per, stride)
        if self._xbuf is None or tuple(self._xbuf[0].shape) != want:
            self._xbuf = [torch.empty(want, dtype=torch.uint8, device=f"cuda:{self.local.device}") for _ in range(2)]
        return self._xbuf

    def _process_views_sharded(self, cams, masks, H, W, invert):
        """Slab mode, host masks (all ``H x W`` uint8): chunk c of the views is uploaded 1/G per rank (carving:
        and bit-packed), all-gathered over NVLink and carved / averaged, while chunk c+1 is being uploaded —
        two exchange buffers, ordering by CUDA events between torch's stream (copies, NCCL) and the engine's
        stream (pack, carve / average); no host synchronisation inside the loop.  Chunks are contiguous view
        ranges in the caller's order, so the averaging sums keep the sequential view order (bit-identical)."""
        import torch
        dist = _dist()
        n, G, r = len(masks), self.world, self.rank
        carving = self.type == "carving"
        stride = self.local.packed_size(H, W) if carving else ((H * W + 255) // 256) * 256
        cams = np.ascontiguousarray(cams, dtype=np.float32).reshape(n, 16)
        spec = carving and bool(getattr(self.local, "speculative_readback", False)) and hasattr(self.local, "spec_snapshot")
        if carving:
            # carving is order-independent: visit the views in farthest-point order of their optical axes (the first
            # few remove almost all empty space).  After the first chunk (32 views) the volume as it stands starts its
            # way to the host (speculative read-back) under the upload / exchange / carve of the rest, which goes in
            # chunks of ~CARVE_CHUNK_VIEWS: the host blocks in the upload of chunk c+1 while the all-gather and the
            # kernels of chunk c run (one chunk for the rest left upload -> all-gather -> kernels in a row: 44 ms
            # instead of 33 at G = 2, tools/e2e_trace.py)
            order = spread_order(cams)
            cams = cams[order]
            masks = [masks[v] for v in order]
            bounds = carve_chunk_bounds(n, G, int(os.environ.get("BP_SHARD_CHUNK_VIEWS", self.CARVE_CHUNK_VIEWS)))
        else:
            # averaging: ~CHUNK_BYTES of maps per chunk, upload hidden behind the kernel; the caller's view order is kept
            per_chunk = max(1, self.CHUNK_BYTES // (H * W))
            per_chunk = min(n, max(G, (per_chunk // G) * G))  # a multiple of G, at most n
            bounds = list(range(0, n, per_chunk)) + [n]
        n_per = max((b - a + G - 1) // G for a, b in zip(bounds[:-1], bounds[1:]))
        bufs = self._exchange_buffers(n_per, stride)
        dev = bufs[0].device
        cur = torch.cuda.current_stream(dev)
        eng = torch.cuda.ExternalStream(self.local.stream(), device=dev)
        done = [None, None]  # event after the pass that last read each buffer
        self._tick("start")
        for ci, (c0, c1) in enumerate(zip(bounds[:-1], bounds[1:])):
            buf = bufs[ci & 1]
            mine = list(range(c0 + r, c1, G))
            hw = np.tile(np.array([[H, W]], dtype=np.int32), (c1 - c0, 1))
            if done[ci & 1] is not None:
                cur.wait_event(done[ci & 1])      # NCCL / copies must not overwrite what a pass still reads
            if carving:
                # (the engine's own stream is already ordered behind that pass)
                # with the cameras the library uploads and packs only the rectangle of each view that the whole grid
                # projects into (BP_VIEW_RECT=0 switches that off)
                rect = os.environ.get("BP_VIEW_RECT", "1") not in ("", "0")
                self.local.pack_views([masks[v] for v in mine], buf[r].data_ptr(), stride, invert=invert,
                                      **({"cams": cams[mine]} if rect else {}))
                ev = torch.cuda.Event()
                ev.record(eng)
                cur.wait_event(ev)
            else:
                for q, v in enumerate(mine):
                    buf[r, q, :H * W].view(H, W).copy_(torch.from_numpy(np.ascontiguousarray(masks[v])),
                                                       non_blocking=True)
                self.h2d_torch_bytes += len(mine) * H * W  # uploads the library's own counter does not see
            self._tick(f"chunk {ci}: upload{' + pack' if carving else ''} of {len(mine)} views")
            dist.all_gather_into_tensor(buf.view(-1), buf[r].reshape(-1), group=self.group)
            self._tick(f"chunk {ci}: all-gather")
            ev = torch.cuda.Event()
            ev.record(cur)
            eng.wait_event(ev)
            base = buf.data_ptr()
            ptrs = [base + (((v - c0) % G) * n_per + (v - c0) // G) * stride for v in range(c0, c1)]
            if carving:
                self.local.process_packed_views(cams[c0:c1], ptrs, hw)
                if spec and ci == 0 and len(bounds) > 2:
                    self.local.spec_snapshot()
            else:
                self.local.process_device_views(cams[c0:c1], ptrs, hw, invert=invert)
            self._tick(f"chunk {ci}: kernels on {c1 - c0} views" + (" + early copy of the volume" if spec and ci == 0 and len(bounds) > 2 else ""))
            done[ci & 1] = torch.cuda.Event()
            done[ci & 1].record(eng)
        for ev in done:                            # host masks are not retained past the call
            if ev is not None:
                ev.synchronize()

    def _before_more_views(self):
        """The reference API allows process_view, get_values, process_view, get_values (cl.py:190-232).  Under view
        sharding ``get_values`` leaves the all-reduced volume on every rank: for averaging (SUM) only rank 0 may keep it
        as the base of further accumulation, the others restart from 0 (their seed).  MIN (carving) is idempotent."""
        if self._holds_total:
            if self.rank != 0:
                self.local.clear()
            self._holds_total = False

    def process_device_views(self, cams, mask_ptrs, hw, invert=False):
        """Masks resident in HBM on every rank (device addresses of uint8 / float32 arrays): nothing to exchange,
        every rank ingests from its own copy (packing 1/G each and all-gathering the tiles was measured slower
        than packing everything locally at 3.7 TB/s: profiles/r01_bench_n2_*.json)."""
        self._before_more_views()
        if self.sharding == "views":
            mine = self.my_views(len(mask_ptrs))
            cams = np.ascontiguousarray(cams, dtype=np.float32).reshape(-1, 16)[mine]
            hw = np.ascontiguousarray(hw, dtype=np.int32).reshape(-1, 2)[mine]
            mask_ptrs = [mask_ptrs[v] for v in mine]
        if len(mask_ptrs):
            self.local.process_device_views(cams, mask_ptrs, hw, invert=invert)
        self._reduced = False

    def process_views(self, views, invert=False):
        """``views``: the FULL sequence of ``(intrinsics, rot, tvec, mask)``, identical on every rank
        (a rank only touches the masks of the views it owns)."""
        views = list(views)
        self._before_more_views()
        if views and self._can_shard_ingest([np.shape(v[3]) for v in views], [np.asarray(v[3]).dtype for v in views]):
            n = len(views)
            cams = np.empty((n, 16), dtype=np.float32)
            for q, (intr, rot, tvec, _) in enumerate(views):
                cams[q, 0:4] = np.asarray(intr, dtype=np.float32).reshape(4)
                cams[q, 4:13] = np.asarray(rot, dtype=np.float32).reshape(9)
                cams[q, 13:16] = np.asarray(tvec, dtype=np.float32).reshape(3)
            H, W = np.shape(views[0][3])
            self._process_views_sharded(cams, [v[3] for v in views], int(H), int(W), invert)
            self._reduced = False
            self.last_ingest = "sharded"
            return
        self.last_ingest = "replicated"
        mine = [views[v] for v in self.my_views(len(views))]
        self.local.process_views(mine, invert=invert)
        self._reduced = False

    def process_view(self, intrinsics, rot, tvec, mask):
        """One view on every rank's grid part (slabs) / on the rank whose turn it is (views)."""
        self._before_more_views()
        if self.sharding == "views":
            turn = getattr(self, "_turn", 0)
            self._turn = turn + 1
            if turn % self.world != self.rank:
                self._reduced = False
                return
        self.local.process_view(intrinsics, rot, tvec, mask)
        self._reduced = False

    def process_label(self, fs, camera_metadata, label=None, invert=False, gather=True):
        """cl.py:259-305 distributed: the file loop (channel filter, camera lookup, warn + skip) runs
        on every rank so that all agree on the view list; only owned views are decoded."""
        if hasattr(fs, "get_files"):
            fs = fs.get_files()
        usable = []
        for fi in fs:
            if label is not None and fi.get_metadata("channel") != label:
                continue
            cam = fi.get_metadata(camera_metadata, default=None)
            if cam is None:
                logger.warning(f"Could not get camera params from '{camera_metadata}' for {fi.id}, skipping...")
                continue
            usable.append((fi, cam))
        from .cl import _read_image
        self._before_more_views()
        mine = []
        for v in self.my_views(len(usable)):
            fi, cam = usable[v]
            intrinsics = np.array(cam["camera_model"]["params"][0:4], dtype=np.float32)
            rot = np.array(sum(cam["rotmat"], []), dtype=np.float32)
            tvec = np.array(cam["tvec"], dtype=np.float32)
            mine.append((intrinsics, rot, tvec, _read_image(fi)))
        self.local.process_views(mine, invert=invert)
        self._reduced = False
        return self.get_values(gather=gather)

    def process_fileset(self, fs, camera_metadata, invert=False, gather=True, exp_clip=False, stats=None):
        """cl.py:234-257 distributed (float64 ``[L, ...]`` stack when labels are set).  ``exp_clip`` / ``stats``: the
        same contract as ``Backprojection.process_fileset`` (the tail of ``Voxels.run``, tasks/cl.py:166-174); here
        the min / max are combined over the ranks and ``np.exp`` + clip run on the host after the exchange."""
        if self.labels is not None:
            result = None
            for i, label in enumerate(self.labels):
                if i != 0:
                    self.clear()
                vol = self.process_label(fs, camera_metadata, label, invert, gather=gather)
                if result is None:
                    result = np.zeros((len(self.labels), *vol.shape))
                result[i, :] = vol
        else:
            result = self.process_label(fs, camera_metadata, None, invert=invert, gather=gather)
        if stats is not None:
            lo = float(result.min()) if result.size else np.inf
            hi = float(result.max()) if result.size else -np.inf
            if self.world > 1 and not (gather or self.sharding == "views"):
                import torch
                dist = _dist()
                nccl = str(dist.get_backend(self.group)) == "nccl"
                dev = torch.device("cuda", int(getattr(self.local, "device", 0))) if nccl else torch.device("cpu")
                t = torch.tensor([-lo, hi], dtype=torch.float64, device=dev)  # NaN propagates through MAX
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
                lo, hi = -float(t[0]), float(t[1])
            stats["min"], stats["max"] = lo, hi
        if exp_clip:
            if self.type != "averaging":
                raise ValueError("exp/clip applies to averaging volumes only (tasks/cl.py:172)")
            if self.labels is None:
                result = result.copy()  # get_values aliases a reused host buffer
            np.exp(result, out=result)
            result[result > 1] = 1.0
        return result

    # -- combination ---------------------------------------------------------------------------
    def reduce(self):
        """View sharding: all-reduce the raw state in place (MIN int8 / SUM float32).  Idempotent
        until more views are processed.  Slab sharding: nothing to exchange."""
        if self.sharding != "views" or self.world == 1 or self._reduced:
            self._reduced = True
            return
        dist = _dist()
        state = self.local.state_tensor()
        self.local.sync()
        op = dist.ReduceOp.MIN if self.type == "carving" else dist.ReduceOp.SUM
        dist.all_reduce(state, op=op, group=self.group)
        self.local.adopt_state(state)
        self._reduced = True
        self._holds_total = self.type == "averaging"

    def _planes_of(self, rank):
        nx = self.shape[0]
        if self.sharding == "slabs":
            a, b = slab_range(nx, rank, self.world)
            return np.arange(a, b)
        lg = self._stripe_lg
        return np.array([i for i in range(nx) if (i >> lg) % self.world == rank], dtype=np.int64)

    def _gather_planes(self, vol):
        """Plane sharding: every rank's ``[n_local, ny, nz]`` part -> the full ``[nx, ny, nz]`` volume on every rank.
        ONE tensor all-gather (the parts padded to the largest plane count).  With NCCL and a CUDA engine the raw
        device state travels over NVLink - int8 for carving, a quarter of the reference's int32 - and is decoded on
        the host after a single D2H; on CPU ranks (gloo) the numpy parts are gathered."""
        import torch
        dist = _dist()
        counts = [len(self._planes_of(r)) for r in range(self.world)]
        n_max = max(counts)
        nccl = str(dist.get_backend(self.group)) == "nccl"
        if nccl and hasattr(self.local, "device_state"):
            self.local.sync()
            st = self.local.state_tensor()  # [n_local, ny, nz_pitch] int8 / float32, aliases the engine's state
            send = torch.zeros((n_max,) + tuple(st.shape[1:]), dtype=st.dtype, device=st.device)
            send[:st.shape[0]].copy_(st)
            recv = torch.empty((self.world,) + tuple(send.shape), dtype=st.dtype, device=st.device)
            dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=self.group)
            parts = recv.cpu().numpy()[..., :self.shape[2]]
            if self.type == "carving":
                parts = decode_carve_state(parts, self.default_value)
        else:
            send = torch.zeros((n_max,) + tuple(vol.shape[1:]), dtype=torch.from_numpy(np.empty(0, vol.dtype)).dtype)
            send[:vol.shape[0]] = torch.from_numpy(np.ascontiguousarray(vol))
            recv = [torch.empty_like(send) for _ in range(self.world)]
            dist.all_gather(recv, send, group=self.group)
            parts = [t.numpy() for t in recv]
        out = np.empty((self.shape[0], *vol.shape[1:]), dtype=vol.dtype)
        for r in range(self.world):
            out[self._planes_of(r)] = parts[r][:counts[r]]
        return out

    def get_values(self, gather=True):
        """The volume (cl.py:229-232).  Views: the full grid on every rank, after ``reduce()``.
        Slabs: this rank's planes, or with ``gather`` the full grid all-gathered on every rank."""
        self.reduce()
        if self.sharding == "views" or self.world == 1 or not gather:
            return self.local.get_values()
        if str(_dist().get_backend(self.group)) == "nccl" and hasattr(self.local, "device_state"):
            return self._gather_planes(np.empty((0, self.shape[1], self.shape[2]), dtype=self.dtype))
        return self._gather_planes(self.local.get_values())

    get_volume = get_values

    def clear(self):
        self.local.clear()
        self._turn = 0
        self._reduced = True
        self._holds_total = False

    def close(self):
        self.local.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

"""Worker side of tests/test_parallel_gloo.py: CPU ranks (gloo) running the PRODUCT's sharding logic
(plant3dvision_b200.parallel) over a stand-in engine backed by the CPU oracle.  Test infrastructure."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class OracleEngine:
    """Duck-types what ``DistributedBackprojection`` needs from ``cl.Backprojection``."""

    def __init__(self, shape, origin, voxel_size, type="carving", default_value=0, labels=None, log=False,
                 i_range=None, stripes=None):
        from oracle.backprojection_oracle import BackprojectionOracle
        # stripes = (lg_block, world, rank): the stand-in computes the whole grid and hands out its planes
        self.planes = None
        if stripes is not None:
            lg, world, rank = stripes
            self.planes = np.array([i for i in range(shape[0]) if (i >> lg) % world == rank])
        self.i_range = i_range or (0, shape[0])
        self.default_value = default_value
        self.type = type
        local = [self.i_range[1] - self.i_range[0], shape[1], shape[2]]
        self.o = BackprojectionOracle(local, origin, voxel_size, type=type, default_value=default_value,
                                      labels=labels, log=log, i_offset=self.i_range[0])

    def process_views(self, views, invert=False):
        for (i, r, t, m) in views:
            self.o.process_view(i, r, t, np.invert(m) if invert else m)

    def process_view(self, i, r, t, m):
        self.o.process_view(i, r, t, m)

    def get_values(self):
        vol = self.o.get_values()
        return vol if self.planes is None else vol[self.planes]

    def clear(self):
        self.o.clear()

    def close(self):
        pass

    def sync(self):
        pass

    def state_tensor(self):
        import torch
        from plant3dvision_b200.parallel import encode_carve_state
        if self.type == "carving":
            return torch.from_numpy(encode_carve_state(self.o.get_values(), self.default_value))
        return torch.from_numpy(self.o.get_values())  # aliases the oracle's buffer

    def adopt_state(self, state):
        from plant3dvision_b200.parallel import decode_carve_state
        if self.type == "carving":
            self.o.values_h = decode_carve_state(state.numpy(), self.default_value)


def scene(seed=5, n_views=7, H=120, W=160):
    rng = np.random.default_rng(seed)
    shape, vs, origin = [21, 17, 30], 1.5, [-15.0, -12.0, -20.0]
    views = []
    for v in range(n_views):
        th = 2 * np.pi * v / n_views
        C = np.array([90 * np.cos(th), 90 * np.sin(th), 10.0])
        fwd = -C / np.linalg.norm(C)
        right = np.cross(fwd, [0, 0, 1.0])
        right /= np.linalg.norm(right)
        R = np.stack([right, np.cross(fwd, right), fwd])
        intr = np.array([150.0, 150.0, W / 2, H / 2], dtype=np.float32)
        yy, xx = np.mgrid[0:H, 0:W]
        blob = ((xx - W / 2) ** 2 / (30 + 5 * v) ** 2 + (yy - H / 2) ** 2 / 45 ** 2) < 1
        mask = (blob * rng.integers(1, 256, (H, W))).astype(np.uint8)
        views.append((intr, R.astype(np.float32).reshape(-1), (-R @ C).astype(np.float32), mask))
    return shape, origin, vs, views


def worker(rank, world, port, kind, sharding, default_value, out_dir, backend="gloo"):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    engine_kwargs = {"engine": OracleEngine}
    if backend == "nccl":  # the real thing: one GPU per rank, CUDA engine, NCCL all-reduce of device state
        torch.cuda.set_device(rank)
        engine_kwargs = {"device": rank}
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from plant3dvision_b200.parallel import DistributedBackprojection
        shape, origin, vs, views = scene()
        if backend == "nccl" and sharding in ("slabs", "stripes"):
            engine_kwargs["ingest"] = "sharded"  # exercise the NVLink exchange for both kernels at world size 2
        bp = DistributedBackprojection(shape, origin, vs, type=kind, default_value=default_value, log=False,
                                       sharding=sharding, **engine_kwargs)
        bp.process_views(views)
        vol = np.array(bp.get_values(gather=True))  # copy: get_values aliases a reused host buffer
        again = bp.get_values(gather=True)  # reduce() must be idempotent
        assert np.array_equal(vol, again)
        part = bp.get_values(gather=False)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), vol=vol, part=part, i_range=np.array(bp.i_range))
        bp.clear()
        for v in views:  # per-view entry point, every rank calling it for every view
            bp.process_view(*v)
        assert np.array_equal(bp.get_values(gather=True), vol) or kind == "averaging"
        if kind == "averaging":
            np.testing.assert_allclose(bp.get_values(gather=True), vol, rtol=1e-5, atol=1e-6)
        if backend == "nccl" and sharding in ("slabs", "stripes"):
            # masks resident in HBM on every rank (bench.py's `value` leg): sharded pack + all-gather of bit tiles
            bp.clear()
            dev = [torch.from_numpy(v[3]).cuda(rank) for v in views]
            cams = np.stack([np.concatenate([np.ravel(v[0]), np.ravel(v[1]), np.ravel(v[2])]) for v in views])
            hw = np.array([m.shape for m in dev], dtype=np.int32)
            bp.process_device_views(cams.astype(np.float32), [m.data_ptr() for m in dev], hw)
            assert np.array_equal(bp.get_values(gather=True), vol)
            # the chunked pipeline of the sharded ingest (2 views per chunk -> 4 chunks, both exchange buffers reused)
            bp.clear()
            type(bp).CHUNK_BYTES = 3 * views[0][3].size
            bp._xbuf = None
            bp.process_views(views)
            assert np.array_equal(bp.get_values(gather=True), vol)
            if kind == "carving":
                # more views than the head chunk of the sharded ingest: the volume starts its way to the host after
                # the first chunk (speculative read-back across ranks), get_values(gather=False) patches what changed
                from oracle.backprojection_oracle import BackprojectionOracle
                shape40, origin40, vs40, views40 = scene(seed=9, n_views=40)
                bp40 = DistributedBackprojection(shape40, origin40, vs40, type=kind, default_value=default_value,
                                                 sharding=sharding, **engine_kwargs)
                ora = BackprojectionOracle(shape40, origin40, vs40, type=kind, default_value=default_value)
                for v in views40:
                    ora.process_view(*v)
                for rep in range(2):
                    bp40.clear()
                    bp40.process_views(views40)
                    part40 = np.array(bp40.get_values(gather=False))
                    assert np.array_equal(part40, ora.get_values()[bp40.planes]), f"rep {rep}"
                assert bp40.local.spec_stats()[1] > 0  # the speculative path did run
                bp40.close()
            # and the inverted form against the host-side inversion
            bp.clear()
            bp.process_views(views, invert=True)
            a = np.array(bp.get_values(gather=True))
            bp.clear()
            bp.process_views([(i, r, t, np.invert(m)) for (i, r, t, m) in views])
            assert np.array_equal(bp.get_values(gather=True), a)
    finally:
        dist.destroy_process_group()


def worker_incremental(rank, world, port, kind, sharding, out_dir):
    """The reference API allows process_view(s), get_values, process_view(s), get_values (cl.py:190-232): the second
    result must be the sequential one over ALL views, not the first total counted once per rank."""
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from plant3dvision_b200.parallel import DistributedBackprojection
        shape, origin, vs, views = scene()
        bp = DistributedBackprojection(shape, origin, vs, type=kind, default_value=0.25 if kind == "averaging" else 0,
                                       log=False, sharding=sharding, engine=OracleEngine)
        bp.process_views(views[:3])
        first = np.array(bp.get_values(gather=True))
        bp.process_views(views[3:5])
        second = np.array(bp.get_values(gather=True))
        for v in views[5:]:
            bp.process_view(*v)
        third = np.array(bp.get_values(gather=True))
        np.savez(os.path.join(out_dir, f"inc{rank}.npz"), first=first, second=second, third=third)
    finally:
        dist.destroy_process_group()


def worker_voxels_task(rank, world, port, scan_dir, sharding, out_dir):
    """The Voxels task mirror over a distributed engine: every rank computes, rank 0 writes."""
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from plant3dvision_b200.parallel import DistributedBackprojection
        from plant3dvision_b200.tasks import cl as tcl
        out = tcl.run_voxels(scan_dir, {"type": "averaging", "voxel_size": 1.0, "log": True, "camera_metadata": "camera",
                                        "upstream_mask": "ImagesFilesetExists", "labels": ["stem", "background"]},
                             backprojection_class=DistributedBackprojection,
                             bp_kwargs={"engine": OracleEngine, "sharding": sharding})
        assert (out is None) == (rank != 0)
        # the distributed tail contract on its own: local planes, min / max agreed over the ranks
        from plant3dvision_b200 import fsdb_lite
        files = fsdb_lite.Scan(scan_dir).get_fileset("images").get_files()
        bp = DistributedBackprojection([9, 7, 5], [-4.0, -3.0, -2.0], 1.0, type="averaging", labels=["stem"], log=True,
                                       sharding="slabs", engine=OracleEngine)
        st = {}
        part = bp.process_fileset(files, "camera", gather=False, exp_clip=True, stats=st)
        np.savez(os.path.join(out_dir, f"tail{rank}.npz"), part=part, lo=st["min"], hi=st["max"], i_range=np.array(bp.i_range))
    finally:
        dist.destroy_process_group()

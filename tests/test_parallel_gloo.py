"""world_size-2 ``gloo`` tests (CPU) of the multi-GPU host logic in ``plant3dvision_b200.parallel``:
partitioning, the monotone carve-state coding under an all-reduce MIN, the SUM for averaging, and
slab sharding + gather.  The per-rank compute is the CPU oracle standing in for the CUDA engine;
the expected result is the oracle's sequential single-process run over all views."""
import socket

import numpy as np
import pytest

from _gloo_worker import scene, worker
from oracle.backprojection_oracle import BackprojectionOracle
from plant3dvision_b200 import parallel


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def spawn(fn, world, *args):
    """mp.spawn(fn, (world, port, *args)); a rendezvous that lost its port to another process between free_port()
    and the bind (or hit a lingering socket) is retried on a fresh port - anything else fails at once."""
    import torch.multiprocessing as mp
    for attempt in range(3):
        try:
            return mp.spawn(fn, args=(world, free_port(), *args), nprocs=world, join=True)
        except Exception as e:  # noqa: BLE001 - the child's traceback arrives as text
            text = str(e).lower()
            if attempt == 2 or not any(k in text for k in ("address already in use", "eaddrinuse", "connection refused",
                                                           "connection reset", "timed out", "timeout")):
                raise


def sequential(kind, default_value):
    shape, origin, vs, views = scene()
    o = BackprojectionOracle(shape, origin, vs, type=kind, default_value=default_value)
    for v in views:
        o.process_view(*v)
    return o.get_values()


@pytest.mark.parametrize("kind,sharding,default_value", [
    ("carving", "views", 0), ("carving", "slabs", 0), ("carving", "views", 3),
    ("averaging", "views", 0), ("averaging", "slabs", 0), ("averaging", "views", 0.5),
    ("carving", "stripes", 0), ("averaging", "stripes", 0),
])
def test_two_ranks_equal_sequential(tmp_path, kind, sharding, default_value):
    world = 2
    spawn(worker, world, kind, sharding, default_value, str(tmp_path))
    want = sequential(kind, default_value)
    for rank in range(world):
        z = np.load(tmp_path / f"rank{rank}.npz")
        if kind == "carving" or sharding in ("slabs", "stripes"):
            assert np.array_equal(z["vol"], want)  # bit-exact
        else:
            np.testing.assert_allclose(z["vol"], want, rtol=1e-5, atol=1e-6)  # summation order changed
        i0, i1 = z["i_range"]
        if sharding == "stripes":  # blocks of 8 planes dealt round robin (21 planes: rank 0 holds 0-7 and 16-20)
            planes = [i for i in range(want.shape[0]) if (i >> 3) % world == rank]
            assert np.array_equal(z["part"], want[planes])
            continue
        assert (i0, i1) == (parallel.slab_range(want.shape[0], rank, world) if sharding == "slabs" else (0, want.shape[0]))
        if sharding == "slabs":
            assert np.array_equal(z["part"], want[i0:i1])


@pytest.mark.parametrize("kind,sharding", [("averaging", "views"), ("carving", "views"), ("averaging", "stripes")])
def test_more_views_after_get_values(tmp_path, kind, sharding):
    """under view sharding a SUM all-reduce leaves the total on every rank; views processed after a
    get_values() must be added to it ONCE, not to world-size copies of it."""
    from _gloo_worker import worker_incremental
    world = 2
    spawn(worker_incremental, world, kind, sharding, str(tmp_path))
    shape, origin, vs, views = scene()
    o = BackprojectionOracle(shape, origin, vs, type=kind, default_value=0.25 if kind == "averaging" else 0)
    want = []
    for upto in (3, 5, len(views)):
        while len(want) < 3 and getattr(o, "_n", 0) < upto:
            o.process_view(*views[getattr(o, "_n", 0)])
            o._n = getattr(o, "_n", 0) + 1
        want.append(o.get_values().copy())
    for rank in range(world):
        z = np.load(tmp_path / f"inc{rank}.npz")
        for got, w in zip((z["first"], z["second"], z["third"]), want):
            if kind == "carving" or sharding != "views":
                assert np.array_equal(got, w)
            else:
                np.testing.assert_allclose(got, w, rtol=1e-5, atol=1e-6)


def test_slab_range_and_view_partition_cover_everything():
    for nx in (1, 7, 128, 1024, 1025):
        for world in (1, 2, 3, 8):
            edges = [parallel.slab_range(nx, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == nx
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            assert max(e[1] - e[0] for e in edges) - min(e[1] - e[0] for e in edges) <= 1
    for n in (0, 1, 5, 240):
        for world in (1, 2, 4, 8):
            got = sorted(sum([parallel.partition_views(n, r, world) for r in range(world)], []))
            assert got == list(range(n))


def test_carve_state_coding_is_monotone_and_round_trips():
    rng = np.random.default_rng(0)
    a = rng.choice([-1, 0, 1], size=(4, 5, 6)).astype(np.int32)
    b = rng.choice([-1, 0, 1], size=(4, 5, 6)).astype(np.int32)
    assert np.array_equal(parallel.decode_carve_state(parallel.encode_carve_state(a)), a)
    # sequential semantics of two replicas: carved if either carved, else seen if either seen
    want = np.where((a == -1) | (b == -1), -1, np.where((a == 1) | (b == 1), 1, 0))
    got = parallel.decode_carve_state(np.minimum(parallel.encode_carve_state(a), parallel.encode_carve_state(b)))
    assert np.array_equal(got, want)
    assert not np.array_equal(np.minimum(a, b), want)  # why the raw {-1,0,1} coding cannot be MIN-reduced
    # non-zero seed: `seen` is not representable (the kernel only promotes == 0), -1 seed is inert
    c = np.where(a == -1, -1, 7).astype(np.int32)
    assert np.array_equal(parallel.decode_carve_state(parallel.encode_carve_state(c, 7), 7), c)
    assert (parallel.decode_carve_state(parallel.encode_carve_state(a), -1) == -1).all()


def test_bad_arguments():
    with pytest.raises(ValueError):
        parallel.DistributedBackprojection([2, 2, 2], [0, 0, 0], 1.0, sharding="nope")
    with pytest.raises(ValueError):
        parallel.DistributedBackprojection([2, 2, 2], [0, 0, 0], 1.0, type="nope")


@pytest.mark.parametrize("sharding", ["views", "stripes"])
def test_voxels_task_over_a_distributed_engine(tmp_path, sharding):
    """`run_voxels(..., backprojection_class=DistributedBackprojection)` on two ranks: rank 0 writes the same
    Voxels.npz a single process writes (tasks/cl.py:149-186), the tail (exp + clip, min / max) included."""
    from _gloo_worker import worker_voxels_task
    from test_host_logic import make_scan
    from test_tasks_cpu import OracleEngine as SingleEngine
    from plant3dvision_b200 import fsdb_lite
    scan_dir = str(make_scan(tmp_path, n=4, channels=("background", "stem")))
    world = 2
    spawn(worker_voxels_task, world, scan_dir, sharding, str(tmp_path))
    scan = fsdb_lite.Scan(scan_dir)
    outs = [fs for fs in scan.get_filesets() if fs.id.startswith("Voxels_")]
    assert len(outs) == 1 and [f.filename for f in outs[0].get_files()] == ["Voxels.npz"]
    files = scan.get_fileset("images").get_files()
    eng = SingleEngine([9, 7, 5], [-4.0, -3.0, -2.0], 1.0, type="averaging", labels=["stem", "background"], log=True)
    st = {}
    want = eng.process_fileset(files, "camera", exp_clip=True, stats=st)
    with np.load(outs[0].get_files()[0].path()) as z:
        assert z.files == ["stem", "background"]
        for i, l in enumerate(z.files):
            if sharding == "views":  # float32 sums in another order, then exp
                np.testing.assert_allclose(z[l], want[i], rtol=1e-5, atol=1e-7)
            else:
                assert np.array_equal(z[l], want[i])
    # local planes + agreed min / max
    raw = SingleEngine([9, 7, 5], [-4.0, -3.0, -2.0], 1.0, type="averaging", labels=["stem"], log=True)
    st1 = {}
    full = raw.process_fileset(files, "camera", exp_clip=True, stats=st1)
    for rank in range(world):
        z = np.load(tmp_path / f"tail{rank}.npz")
        i0, i1 = z["i_range"]
        assert np.array_equal(z["part"], full[:, i0:i1])
        assert float(z["lo"]) == st1["min"] and float(z["hi"]) == st1["max"]


def test_carve_chunk_bounds_cover_every_view_once():
    """Chunks of the sharded carving ingest (parallel.py): contiguous, increasing, first 0 / last n, the first chunk the
    32-view head (rounded up to the world size), later chunks multiples of the world size except possibly the last,
    no short last chunk."""
    from plant3dvision_b200.parallel import carve_chunk_bounds
    for n in (0, 1, 5, 18, 31, 32, 33, 60, 72, 120, 240, 241, 1000):
        for G in (1, 2, 3, 4, 8):
            for chunk in (1, 7, 48, 64, 1000):
                b = carve_chunk_bounds(n, G, chunk)
                assert b[0] == 0 and b[-1] == n and all(x < y for x, y in zip(b, b[1:])) or (n == 0 and b == [0, 0]), (n, G, chunk, b)
                if n > 0 and len(b) > 2:
                    step = max(G, (chunk // G) * G)
                    assert b[1] == -(-32 // G) * G
                    assert all((y - x) % G == 0 for x, y in zip(b[1:-1], b[2:-1])), (n, G, chunk, b)
                    assert all(y - x in (step, 2 * step) or y == n for x, y in zip(b[1:], b[2:])), (n, G, chunk, b)
                    if len(b) > 3:
                        assert b[-1] - b[-2] >= step // 2, (n, G, chunk, b)
    assert carve_chunk_bounds(240, 8, 48) == [0, 32, 80, 128, 176, 240]
    assert carve_chunk_bounds(18, 2, 48) == [0, 18]

"""CPU tests of the task mirrors' HOST logic (file selection, ordering, metadata, output formats, the fused-tail
contract): the one device call of each task is replaced by its oracle through monkeypatching, so that what is checked
here is everything AROUND the kernels.  The kernels themselves are checked by the `-m gpu` tests.
"""
import json

import numpy as np
import pytest

from test_host_logic import make_scan


class OracleEngine:
    """Stands in for ``plant3dvision_b200.cl.Backprojection`` in ``Voxels.compute``: the oracle plus the two
    keyword arguments the product class adds to ``process_fileset``."""

    def __init__(self, shape, origin, voxel_size, type="carving", labels=None, log=False, **kw):
        from oracle.backprojection_oracle import BackprojectionOracle
        self.o = BackprojectionOracle(shape, origin, voxel_size, type=type, labels=labels, log=log)
        self.closed = False

    def process_fileset(self, fs, camera_metadata, invert=False, exp_clip=False, stats=None):
        vol = self.o.process_fileset(fs, camera_metadata, invert)
        if stats is not None:
            stats["min"], stats["max"] = float(vol.min()), float(vol.max())
        if exp_clip:
            vol = np.exp(vol)
            vol[vol > 1] = 1.0
        return vol

    def close(self):
        self.closed = True


@pytest.mark.parametrize("log", [True, False])
def test_voxels_task_tail_contract_and_npz_output(tmp_path, monkeypatch, caplog, log):
    import plant3dvision_b200.cl as bcl
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200.tasks import cl as tcl
    monkeypatch.setattr(bcl, "Backprojection", OracleEngine)
    scan_dir = make_scan(tmp_path, n=3, channels=("background", "stem"))
    out = tcl.run_voxels(str(scan_dir), {"type": "averaging", "voxel_size": 1.0, "log": log, "camera_metadata": "camera",
                                         "upstream_mask": "ImagesFilesetExists", "labels": ["stem", "background"]})
    assert out.filename == "Voxels.npz"
    files = fsdb_lite.Scan(str(scan_dir)).get_fileset("images").get_files()
    eng = OracleEngine([9, 7, 5], [-4.0, -3.0, -2.0], 1.0, type="averaging", labels=["stem", "background"], log=log)
    want = eng.process_fileset(files, "camera", exp_clip=log)
    with np.load(out.path()) as z:
        assert z.files == ["stem", "background"]  # the parameter's order, tasks/cl.py:149-151
        for i, l in enumerate(z.files):
            assert z[l].dtype == np.float64 and np.array_equal(z[l], want[i])
        if log:
            assert max(z[l].max() for l in z.files) <= 1.0
    assert out.get_metadata("origin") == [-4.0, -3.0, -2.0] and out.get_metadata("voxel_size") == 1.0
    # the sanity warning of tasks/cl.py:168-169 fires on a constant volume (no view has a camera -> nothing processed)
    scan2 = make_scan(tmp_path / "b", with_cam=False)
    import logging
    with caplog.at_level(logging.WARNING):
        tcl.run_voxels(str(scan2), {"type": "carving", "voxel_size": 1.0, "camera_metadata": "camera",
                                    "upstream_mask": "ImagesFilesetExists", "query": {"channel": "background"}})
    assert "something WRONG with the volume" in caplog.text


def test_masks_task_host_logic(tmp_path, monkeypatch):
    from PIL import Image
    from oracle import masks_oracle as mo
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200 import proc2d
    from plant3dvision_b200.tasks import proc2d as tp
    calls = []

    def fake_masks(images, type="linear", parameters=(0, 1, 0), threshold=0.3, dilation=0, *, device=None,
                   keep_on_device=False):
        calls.append([im.shape for im in images])
        assert len({im.shape for im in images}) == 1  # one shape per device call
        seq = proc2d.dilation_sequence(dilation) if np.isscalar(dilation) else tuple(dilation)
        return [mo.masks_filter(im, type, list(parameters), threshold, seq) for im in images]

    monkeypatch.setattr(proc2d, "masks", fake_masks)
    rng = np.random.default_rng(3)
    scan_dir = tmp_path / "scan"
    (scan_dir / "images").mkdir(parents=True)
    (scan_dir / "metadata" / "images").mkdir(parents=True)
    files, imgs = [], {}
    for v in range(5):
        fid = f"{v:05d}_rgb"
        shape = (20, 30, 3) if v != 2 else (24, 30, 3)  # a stray image of another size must not break the batch
        imgs[fid] = rng.integers(0, 256, shape, dtype=np.uint8)
        Image.fromarray(imgs[fid]).save(scan_dir / "images" / f"{fid}.png")
        json.dump({"channel": "rgb"}, open(scan_dir / "metadata" / "images" / f"{fid}.json", "w"))
        files.append({"id": fid, "file": f"{fid}.png"})
    json.dump({"filesets": [{"id": "images", "files": files}]}, open(scan_dir / "files.json", "w"))
    fs = tp.run_masks(str(scan_dir), {"upstream_task": "ImagesFilesetExists", "type": "excess_green", "threshold": 0.0,
                                      "dilation": 1})
    assert sorted(len(c) for c in calls) == [1, 4]
    got = fsdb_lite.Scan(str(scan_dir)).get_fileset(fs.id)
    assert got.id.startswith("Masks_") and [f.id for f in got.get_files()] == sorted(imgs)  # file order kept
    for f in got.get_files():
        want = mo.masks_filter(imgs[f.id], "excess_green", [0, 1, 0], 0.0, mo.disk_sequence(1))
        assert f.filename.endswith(".png") and np.array_equal(f.read_image(), want)
        assert f.get_metadata("Masks") == {"upstream_task": "ImagesFilesetExists", "filter": "excess_green",
                                           "threshold": 0.0, "dilation": 1}  # no linear_coeff, no query (:242-247)
    assert got.get_metadata("task_params")["type"] == "excess_green"
    with pytest.raises(ImportError, match="skimage"):
        proc2d.dilation_sequence(3)  # the decomposition table lives in scikit-image (absent here)
    assert proc2d.dilation_sequence(0) == ()


def test_segmented_point_cloud_task_host_logic(tmp_path, monkeypatch):
    from PIL import Image
    from oracle import vol2pcd_oracle as vo
    from plant3dvision_b200 import proc3d
    from plant3dvision_b200.tasks import proc3d as tp
    seen = {}

    def fake_score(points, views, labels, *, device=None, return_scores=True):
        seen["labels"] = list(labels)
        seen.setdefault("calls", []).append([(l, c is not None, m.shape) for l, c, m in views])
        lab, sc = vo.score_points(points, views, list(labels))
        return lab, sc, 0.0

    monkeypatch.setattr(proc3d, "score_points", fake_score)
    scan_dir = tmp_path / "scan"
    for d in ("PointCloud_ab12", "Segmentation2D_cd34", "metadata/Segmentation2D_cd34"):
        (scan_dir / d).mkdir(parents=True)
    rng = np.random.default_rng(4)
    pts = np.column_stack([rng.uniform(-3, 3, 200), rng.uniform(-2, 2, 200), rng.uniform(-1, 1, 200)])
    tp.write_ply(scan_dir / "PointCloud_ab12" / "PointCloud.ply", pts, np.tile([0.0, 0.0, 1.0], (200, 1)))
    cam = {"camera_model": {"params": [20.0, 20.0, 8.0, 6.0]}, "rotmat": np.eye(3).tolist(), "tvec": [0, 0, 30.0]}
    files = []
    for v in range(3):
        for ch in ("background", "rgb", "stem", "leaf", None):
            fid = f"{v:05d}_{ch}"
            Image.fromarray(rng.integers(0, 256, (12, 16), dtype=np.uint8)).save(scan_dir / "Segmentation2D_cd34" / f"{fid}.png")
            md = {"camera": cam} if v != 1 else {}  # view 1 has no pose: skipped with a warning
            if ch is not None:
                md["channel"] = ch
            json.dump(md, open(scan_dir / "metadata" / "Segmentation2D_cd34" / f"{fid}.json", "w"))
            files.append({"id": fid, "file": f"{fid}.png"})
    json.dump({"filesets": [{"id": "PointCloud_ab12", "files": [{"id": "PointCloud", "file": "PointCloud.ply"}]},
                            {"id": "Segmentation2D_cd34", "files": files}]}, open(scan_dir / "files.json", "w"))
    task = tp.SegmentedPointCloud(str(scan_dir), use_colmap_poses=False)
    out = task.run()
    assert seen["labels"] == ["leaf", "stem"]  # background and rgb removed (tasks/proc3d.py:198-201), sorted
    # only the views that will be scored are decoded and handed over: 2 poses x 2 labels, in file order
    assert seen["calls"] == [[("stem", True, (12, 16)), ("leaf", True, (12, 16))] * 2]
    # chunked scoring (bounded memory) adds up to the same labels
    one_shot = out.get_metadata("labels")
    seen["calls"] = []
    monkeypatch.setattr(tp.SegmentedPointCloud, "CHUNK_BYTES", 12 * 16 * 3)
    assert tp.SegmentedPointCloud(str(scan_dir), use_colmap_poses=False).run().get_metadata("labels") == one_shot
    assert [len(c) for c in seen["calls"]] == [3, 1]
    monkeypatch.setattr(tp.SegmentedPointCloud, "CHUNK_BYTES", 2 << 30)
    labels = out.get_metadata("labels")
    assert len(labels) == 200 and set(labels) <= {"leaf", "stem"}
    got_pts, _, col = tp.read_ply(out.path())
    assert np.array_equal(got_pts, pts)
    want = {"leaf": [0, 255, 255], "stem": [0, 0, 255]}
    assert all((col[i] == want[l]).all() for i, l in enumerate(labels))
    # explicit label order decides ties; colmap poses requested but absent -> every view skipped -> all label 0
    seen["calls"] = []
    out2 = tp.SegmentedPointCloud(str(scan_dir), use_colmap_poses=True, labels=["stem", "leaf"]).run()
    assert seen["calls"] == [] and set(out2.get_metadata("labels")) == {"stem"}  # all-zero scores: first label
    assert out2.id == "SegmentedPointCloud"


def test_voxels_registers_as_a_luigi_task_when_the_framework_is_importable(tmp_path, monkeypatch):
    """with luigi + romitask importable, ``plant3dvision_b200.tasks.cl.Voxels`` is a RomiTask with the
    reference's parameters (tasks/cl.py:79-91) and ``requires()`` (:93-97); ``run()`` reads its inputs through the
    task framework and writes through plantdb's io, like the reference (:99-186).  The framework is absent from this
    image, so minimal stand-ins are put in ``sys.modules`` - what is exercised is this package's side of the glue."""
    import sys
    import types
    import plant3dvision_b200.cl as bcl
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200.tasks import cl as tcl

    class Param:
        def __init__(self, default=None):
            self.default = default

    luigi = types.ModuleType("luigi")
    for n in ("Parameter", "DictParameter", "FloatParameter", "BoolParameter", "ListParameter", "TaskParameter"):
        setattr(luigi, n, type(n, (Param,), {}))

    class RomiTask:
        def __init__(self, **kw):
            for k in dir(type(self)):
                v = getattr(type(self), k)
                if isinstance(v, Param):
                    setattr(self, k, kw.get(k, v.default))

    romitask = types.ModuleType("romitask")
    romitask.RomiTask = RomiTask
    written = {}
    plantdb = types.ModuleType("plantdb")
    plantdb.io = types.ModuleType("plantdb.io")
    plantdb.io.write_npz = lambda f, d: written.update(npz=d)
    plantdb.io.write_volume = lambda f, v: written.update(vol=v)
    for name, mod in (("luigi", luigi), ("romitask", romitask), ("plantdb", plantdb), ("plantdb.io", plantdb.io)):
        monkeypatch.setitem(sys.modules, name, mod)
    monkeypatch.setattr(bcl, "Backprojection", OracleEngine)

    Task = tcl.make_luigi_voxels()
    assert Task.__name__ == "Voxels" and issubclass(Task, RomiTask) and Task.upstream_task is None
    want = {"query": {}, "camera_metadata": "colmap_camera", "voxel_size": 1.0, "type": "carving", "log": True,
            "invert": False, "labels": [], "bounding_box": None}
    for k, v in want.items():
        assert getattr(Task, k).default == v, k  # the reference's defaults

    scan_dir = make_scan(tmp_path, n=3, channels=("background", "stem"))
    images = fsdb_lite.Scan(str(scan_dir)).get_fileset("images")

    class Target:
        def __init__(self, fs):
            self.fs = fs

        def get(self):
            return self.fs

    class Upstream:
        @staticmethod
        def get_task_family():
            return "ImagesFilesetExists"

    class OutFile:
        md = None

        def set_metadata(self, md):
            OutFile.md = md

    t = Task(upstream_mask=Upstream, upstream_colmap=Upstream, type="averaging", log=False, camera_metadata="camera",
             labels=["stem", "background"])
    assert set(t.requires()) == {"masks"}  # no Colmap upstream -> only the masks
    monkeypatch.setattr(Upstream, "__call__", lambda self: None, raising=False)
    t.input = lambda: {"masks": Target(images)}
    t.output_file = lambda: OutFile()
    t.run()
    eng = OracleEngine([9, 7, 5], [-4.0, -3.0, -2.0], 1.0, type="averaging", labels=["stem", "background"], log=False)
    ref = eng.process_fileset(images.get_files(), "camera")
    assert list(written["npz"]) == ["stem", "background"]
    assert np.array_equal(written["npz"]["stem"], ref[0]) and np.array_equal(written["npz"]["background"], ref[1])
    assert OutFile.md == {"voxel_size": 1.0, "origin": [-4.0, -3.0, -2.0]}
    # without the framework (this image) the module-level name stays the plain runner
    assert tcl.Voxels is tcl.VoxelsRunner

"""CPU tests of the host side that mirrors the reference interface: ``[Voxels]`` TOML parameters and
defaults (tasks/cl.py:79-91), bounding-box -> shape/origin (tasks/cl.py:127-147), the FSDB reader, the
mask conversions of cl.py:205-215 / 300-301 (which flags / tables are handed to the C ABI), and the
volume writers.  Nothing here launches a kernel."""
import json
import os

import numpy as np
import pytest

from plant3dvision_b200 import _cabi, fsdb_lite
from plant3dvision_b200 import cl as bcl
from plant3dvision_b200.tasks import cl as tcl

# the [Voxels] sections of the reference's configs, verbatim (configs/geom_pipe_virtual.toml:1-8,
# geom_pipe_real.toml:30-39, ml_pipe_real.toml:38-46, ml_pipe_virtual.toml:19-26)
GEOM_VIRTUAL = '''
[Voxels]
upstream_mask = "ImagesFilesetExists"
upstream_colmap = "DummyTask"
query = "{\\"channel\\":\\"background\\"}"  # use the (inverted) background masks
camera_metadata = "camera"
voxel_size = 0.06  # ~9x smaller than real plants
type = "carving"
invert = true
'''
GEOM_REAL = '''
[Voxels]
upstream_mask = "Masks"
upstream_colmap = "Colmap"
camera_metadata = "colmap_camera"
voxel_size = 0.5
type = "carving"
[Voxels.bounding_box]  # default to None
x = [300, 450]
y = [300, 450]
z = [-175, 100]
'''
ML_REAL = '''
[Voxels]
upstream_mask = "Segmentation2D"
upstream_colmap = "Colmap"
use_colmap_poses = true
voxel_size = 1
type = "averaging"
log = false # default to true
invert = false
labels = "[\\"background\\"]"
'''
ML_VIRTUAL = '''
[Voxels]
upstream_mask = "Segmentation2D"
use_colmap_poses = false
voxel_size = 0.3
type = "averaging"
log = false # default to true
invert = false
labels = "[\\"background\\", \\"fruit\\", \\"leaf\\", \\"pedicel\\", \\"stem\\"]"
'''


def cfg(tmp_path, text):
    p = tmp_path / "pipe.toml"
    p.write_text(text)
    return tcl.load_config(str(p))


def test_defaults_match_the_reference_task():
    p = tcl.parse_params({})
    assert p == dict(upstream_mask="Masks", upstream_colmap="Colmap", query={}, camera_metadata="colmap_camera",
                     voxel_size=1.0, type="carving", log=True, invert=False, labels=[], bounding_box=None)


def test_reference_configs_parse(tmp_path):
    p = cfg(tmp_path, GEOM_VIRTUAL)
    assert p["query"] == {"channel": "background"} and p["invert"] is True and p["camera_metadata"] == "camera"
    assert p["voxel_size"] == 0.06 and p["type"] == "carving" and p["upstream_mask"] == "ImagesFilesetExists"
    p = cfg(tmp_path, GEOM_REAL)
    assert p["bounding_box"] == {"x": [300, 450], "y": [300, 450], "z": [-175, 100]} and p["voxel_size"] == 0.5
    p = cfg(tmp_path, ML_REAL)  # unknown key `use_colmap_poses` is ignored, as luigi does
    assert p["labels"] == ["background"] and p["log"] is False and p["type"] == "averaging" and p["voxel_size"] == 1.0
    assert "use_colmap_poses" not in p
    p = cfg(tmp_path, ML_VIRTUAL)
    assert p["labels"] == ["background", "fruit", "leaf", "pedicel", "stem"]
    with pytest.raises(TypeError):
        tcl.Voxels("/nonexistent", use_colmap_poses=True)  # direct construction is strict


def test_grid_from_bbox_known_shapes(vp_inputs):
    # virtual_plant bbox at 0.06 -> [386, 397, 1998]; CI variant 0.1 -> [232, 239, 1199]
    assert tcl.grid_from_bbox(vp_inputs["bbox"], 0.06)[0] == [386, 397, 1998]
    assert tcl.grid_from_bbox(vp_inputs["bbox"], 0.1)[0] == [232, 239, 1199]
    assert tcl.grid_from_bbox(vp_inputs["bbox"], 0.3)[0] == [78, 80, 400]
    real = {"x": [300, 450], "y": [300, 450], "z": [-175, 100]}
    assert tcl.grid_from_bbox(real, 0.5) == ([301, 301, 551], [300, 300, -175])
    assert tcl.grid_from_bbox(real, 1.0)[0] == [151, 151, 276]
    shape, origin = tcl.grid_from_bbox(real, 1.0, {"dx": 1.5, "dy": -2, "dz": 10})  # tasks/cl.py:127-140
    assert shape == [151, 151, 276] and origin == [301.5, 298, -165]


def make_scan(tmp_path, with_cam=True, n=3, channels=("background", "stem")):
    from PIL import Image
    scan = tmp_path / "scan_001"
    (scan / "images").mkdir(parents=True)
    (scan / "metadata" / "images").mkdir(parents=True)
    files = []
    rng = np.random.default_rng(0)
    for v in range(n):
        for ch in channels:
            fid = f"{v:05d}_{ch}"
            Image.fromarray((rng.random((12, 16)) < 0.5).astype(np.uint8) * 255).save(scan / "images" / f"{fid}.png")
            md = {"channel": ch, "shot_id": f"{v:05d}"}
            if with_cam and not (v == 1 and ch == "stem"):
                md["camera"] = {"camera_model": {"params": [20.0, 20.0, 8.0, 6.0, 0, 0, 0, 0]},
                                "rotmat": [[1, 0, 0], [0, 1, 0], [0, 0, 1]], "tvec": [0, 0, 30.0 + v]}
            json.dump(md, open(scan / "metadata" / "images" / f"{fid}.json", "w"))
            files.append({"id": fid, "file": f"{fid}.png"})
    json.dump({"filesets": [{"id": "images", "files": files}]}, open(scan / "files.json", "w"))
    json.dump({"bounding_box": {"x": [-4, 4], "y": [-3, 3], "z": [-2, 2]}, "label_names": list(channels)},
              open(scan / "metadata" / "images.json", "w"))
    json.dump({"displacement": {"dx": 0.0, "dy": 0.0, "dz": 0.0}}, open(scan / "metadata" / "metadata.json", "w"))
    return scan


def test_fsdb_lite_reads_the_on_disk_layout(tmp_path):
    scan = fsdb_lite.Scan(make_scan(tmp_path))
    fs = scan.get_fileset("images")
    assert scan.id == "scan_001" and len(fs.get_files()) == 6
    bg = fs.get_files(query={"channel": "background"})  # fsdb.py:2735-2791 filter
    assert [f.id for f in bg] == ["00000_background", "00001_background", "00002_background"]
    assert fs.get_metadata("bounding_box")["x"] == [-4, 4] and fs.get_metadata("nope", default=5) == 5
    assert bg[0].get_metadata("camera")["tvec"] == [0, 0, 30.0]
    assert fs.get_file("00001_stem").get_metadata("camera", default=None) is None
    img = bg[0].read_image()
    assert img.dtype == np.uint8 and img.shape == (12, 16)
    out = scan.get_fileset("Voxels_x", create=True)
    f = out.create_file("Voxels", "npz")
    f.set_metadata({"voxel_size": 1.0, "origin": [0, 0, 0]})
    again = fsdb_lite.Scan(scan.dir)
    assert again.get_fileset("Voxels_x").get_file("Voxels").get_metadata("voxel_size") == 1.0


def test_voxels_resolves_bbox_labels_and_task_id(tmp_path):
    scan = make_scan(tmp_path)
    t = tcl.Voxels(str(scan), upstream_mask="ImagesFilesetExists", upstream_colmap="DummyTask", camera_metadata="camera",
                   query={"channel": "background"})
    assert t.resolve_bounding_box(None) == {"x": [-4, 4], "y": [-3, 3], "z": [-2, 2]}  # images fileset md (:115-118)
    t2 = tcl.Voxels(str(scan), bounding_box={"x": [0, 1], "y": [0, 1], "z": [0, 1]})
    assert t2.resolve_bounding_box(None)["x"] == [0, 1]  # the parameter wins (:107-109)
    assert t.task_id.startswith("Voxels_") and t.task_id != t2.task_id


def bp_shell(mode, log=False):
    """A Backprojection without a device handle: enough for the pure host conversions."""
    bp = bcl.Backprojection.__new__(bcl.Backprojection)
    bp._mode, bp.log, bp._lut_cache = mode, log, {}
    return bp


def test_img_as_float32_and_tables():
    u8 = np.arange(256, dtype=np.uint8)
    f = bcl.img_as_float32(u8)
    assert f.dtype == np.float32 and f[0] == 0 and f[255] == 1 and f[51] == np.float32(51) * np.float32(1.0 / 255)  # float32 product, as skimage computes it
    assert np.array_equal(bcl.img_as_float32(np.array([True, False])), np.array([1, 0], dtype=np.float32))
    assert bcl.img_as_float32(np.array([65535], dtype=np.uint16))[0] == 1
    bp = bp_shell(_cabi.MODE_AVERAGING, log=True)
    lut = bp._lut("u8")
    assert np.array_equal(lut, np.log(np.float32(1e-10) + f).astype(np.float32)) or np.allclose(lut, np.log(1e-10 + f))
    assert lut.dtype == np.float32 and lut[0] == np.float32(np.log(np.float32(0) + 1e-10))
    assert bp._lut("u8") is lut  # cached: process_views groups views by table identity


@pytest.mark.parametrize("invert", [False, True])
def test_prepare_chooses_dtype_flags_and_tables(invert):
    carve, avg = bp_shell(_cabi.MODE_CARVING), bp_shell(_cabi.MODE_AVERAGING)
    m8 = np.zeros((3, 4), dtype=np.uint8)
    arr, code, flags, lut = carve._prepare(m8, invert)
    assert (code, flags, lut) == (_cabi.MASK_U8, _cabi.FLAG_INVERT if invert else 0, None) and arr is m8
    arr, code, flags, lut = carve._prepare(m8.astype(bool), invert)
    assert (code, flags) == (_cabi.MASK_U8, _cabi.FLAG_BOOL_INVERT if invert else 0) and arr.dtype == np.uint8
    arr, code, flags, lut = carve._prepare(np.full((3, 4), 7, dtype=np.int64), invert)
    assert code == _cabi.MASK_I32 and flags == 0 and (arr == (~7 if invert else 7)).all()  # host np.invert
    arr, code, flags, lut = avg._prepare(m8, invert)
    assert code == _cabi.MASK_U8 and lut is not None and lut.shape == (256,)
    arr, code, flags, lut = avg._prepare(np.full((3, 4), 65535, dtype=np.uint16), invert)
    assert code == _cabi.MASK_F32 and lut is None and arr.dtype == np.float32 and (arr == (0 if invert else 1)).all()
    if invert:
        with pytest.raises(TypeError):
            carve._prepare(np.zeros((3, 4), dtype=np.float32), True)  # np.invert on floats raises (cl.py:300-301)
    with pytest.raises(ValueError):
        carve._prepare(np.zeros((3, 4, 3), dtype=np.uint8), invert)


def test_volume_writers_round_trip(tmp_path):
    vol = (np.arange(4 * 5 * 6).reshape(4, 5, 6) % 3 - 1).astype(np.int32)
    p = str(tmp_path / "v.tiff")
    tcl.write_volume(p, vol)
    back = tcl.read_volume(p)
    assert back.shape == vol.shape and np.array_equal(back, vol)
    volf = np.linspace(0, 1, 4 * 5 * 6, dtype=np.float32).reshape(4, 5, 6)
    tcl.write_volume(p, volf)
    assert np.array_equal(tcl.read_volume(p), volf)


def test_pointcloud_task_defaults_and_ply_round_trip(tmp_path):
    from plant3dvision_b200.tasks import proc3d as tp
    # plant3dvision/tasks/proc3d.py:60-64
    assert tp.DEFAULTS == dict(upstream_task="Voxels", level_set_value=1.0, background_prior=1.0,
                               min_contrast=10.0, min_score=0.2)
    p = tmp_path / "pipe.toml"
    p.write_text('[PointCloud]\nupstream_task = "Voxels"\nlevel_set_value = 1\nunknown_key = 3\n')
    assert tp.load_config(str(p))["level_set_value"] == 1.0
    with pytest.raises(TypeError):
        tp.parse_params({"nope": 1})
    rng = np.random.default_rng(0)
    pts, nrm, col = rng.random((17, 3)), rng.random((17, 3)), rng.integers(0, 2, (17, 3)).astype(float)
    tp.write_ply(tmp_path / "a.ply", pts, nrm, col)
    a, b, c = tp.read_ply(tmp_path / "a.ply")
    assert np.array_equal(a, pts) and np.array_equal(b, nrm) and np.array_equal(c, (col * 255).astype(np.uint8))
    tp.write_ply(tmp_path / "b.ply", pts[:0], nrm[:0])
    a, b, c = tp.read_ply(tmp_path / "b.ply")
    assert a.shape == (0, 3) and c is None
    assert open(tmp_path / "a.ply", "rb").read(3) == b"ply"


# ---------------------------------------------------------------------------------------------
# the NPZ writer of the Voxels task tail
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("force_zip64", [False, True])
def test_parallel_npz_writer_is_read_back_by_numpy(tmp_path, force_zip64):
    import importlib
    import zipfile
    m = importlib.import_module("plant-3d-vision_b200.npz_parallel")
    rng = np.random.default_rng(0)
    arrs = {"background": np.clip(np.exp(rng.normal(-3, 2, (40, 50, 60))), 0, 1),
            "flower": rng.random((50, 60, 7)).astype(np.float32), "empty": np.zeros((0, 3)), "scalar": np.array(3.5),
            "strided": np.arange(10, dtype=np.int32).reshape(2, 5)[:, ::2], "carve": rng.integers(-1, 2, (30, 30, 30)),
            "é": np.ones(3, dtype=bool)}
    p = m.savez_compressed_parallel(tmp_path / "Voxels", arrs, chunk=1 << 14, force_zip64=force_zip64)
    assert p.endswith("Voxels.npz")
    with zipfile.ZipFile(p) as z:
        assert z.testzip() is None
        assert all(i.compress_type == zipfile.ZIP_DEFLATED for i in z.infolist())
        assert [i.filename for i in z.infolist()] == [k + ".npy" for k in arrs]
    with np.load(p) as z:
        assert z.files == list(arrs)
        for k, a in arrs.items():
            assert z[k].dtype == a.dtype and z[k].shape == a.shape and np.array_equal(z[k], a)
    # same members and contents as numpy's own writer
    np.savez_compressed(tmp_path / "ref.npz", **arrs)
    with np.load(tmp_path / "ref.npz") as r, np.load(p) as z:
        assert r.files == z.files and all(r[k].tobytes() == z[k].tobytes() for k in r.files)


def test_parallel_npz_writer_property(tmp_path):
    """Random dtypes / shapes / chunk sizes (chunk boundaries inside and across members, empty and 0-d arrays)."""
    import importlib
    from hypothesis import given, settings, strategies as st
    from hypothesis.extra import numpy as hnp
    m = importlib.import_module("plant-3d-vision_b200.npz_parallel")
    dtypes = st.sampled_from([np.float64, np.float32, np.int32, np.int8, np.uint8, np.bool_, np.int64, np.complex64])
    arrays = st.dictionaries(st.text("abcdefgh_", min_size=1, max_size=6),
                             dtypes.flatmap(lambda d: hnp.arrays(d, hnp.array_shapes(min_dims=0, max_dims=3, min_side=0, max_side=9))),
                             min_size=0, max_size=4)
    counter = [0]

    @settings(max_examples=40, deadline=None)
    @given(arrays, st.sampled_from([1, 7, 64, 4096]), st.booleans(), st.integers(1, 3), st.sampled_from([1, 2, None]))
    def check(arrs, chunk, z64, threads, window):
        counter[0] += 1
        p = m.savez_compressed_parallel(tmp_path / f"a{counter[0]}", arrs, chunk=chunk, force_zip64=z64, threads=threads,
                                        window=window)
        with np.load(p) as z:
            assert sorted(z.files) == sorted(arrs)
            for k, a in arrs.items():
                assert z[k].dtype == a.dtype and z[k].shape == a.shape and z[k].tobytes() == a.tobytes()

    check()


def test_volume_writer_switches_to_bigtiff_for_large_volumes(tmp_path, monkeypatch):
    """A 1024^3 int32 volume does not fit the 32-bit offsets of classic TIFF; the threshold is lowered here so that
    the BigTIFF branch runs on a small array."""
    import struct
    vol = np.random.default_rng(1).integers(-1, 2, (6, 5, 7)).astype(np.int32)
    tcl.write_volume(tmp_path / "classic.tiff", vol)
    monkeypatch.setattr(tcl, "BIGTIFF_BYTES", 100)
    tcl.write_volume(tmp_path / "big.tiff", vol)
    version = lambda p: struct.unpack("<H", open(p, "rb").read(4)[2:4])[0]
    try:
        import tifffile  # noqa: F401  (tifffile chooses by itself; nothing to check)
    except ImportError:
        assert version(tmp_path / "classic.tiff") == 42 and version(tmp_path / "big.tiff") == 43
    for name in ("classic.tiff", "big.tiff"):
        back = tcl.read_volume(tmp_path / name)
        assert back.dtype == vol.dtype and np.array_equal(back, vol)

"""The reference's geometric pipeline around the hot path, on a plain scan directory and on the GPU end to end:
``Masks`` (tasks/proc2d.py) -> ``Voxels`` (tasks/cl.py) -> ``PointCloud`` (tasks/proc3d.py) ->
``SegmentedPointCloud`` (tasks/proc3d.py), each stage checked against its oracle on the files the previous stage
wrote (SURVEY §8 rows a, f1-f4 chained the way configs/geom_pipe_virtual.toml chains them).
"""
import json

import numpy as np
import pytest

from conftest import CHANNELS, vp_shape_origin

pytestmark = pytest.mark.gpu


def test_masks_voxels_pointcloud_segmentation_chain(tmp_path, vp_inputs):
    from PIL import Image
    from oracle import masks_oracle as mo
    from oracle import vol2pcd_oracle as vo
    from oracle.backprojection_oracle import BackprojectionOracle
    from plant3dvision_b200 import fsdb_lite
    from plant3dvision_b200.tasks import cl as tcl
    from plant3dvision_b200.tasks import proc2d as tp2
    from plant3dvision_b200.tasks import proc3d as tp3

    n_views, vs = 12, 0.6
    scan_dir = tmp_path / "scan"
    for d in ("images", "Segmentation2D_1", "metadata/images", "metadata/Segmentation2D_1"):
        (scan_dir / d).mkdir(parents=True)
    images, seg = [], []
    for v in range(n_views):
        c = vp_inputs["cams"][v]
        cam = {"camera_model": {"params": [float(x) for x in c[0:4]] + [0, 0, 0, 0]},
               "rotmat": c[4:13].reshape(3, 3).tolist(), "tvec": c[13:16].tolist()}
        # an RGB rendering whose green channel is the plant probability (255 - background)
        bg = vp_inputs["maps"]["background"][v]
        rgb = np.stack([bg // 4, 255 - bg, bg // 5], axis=2).astype(np.uint8)
        fid = f"{v:05d}_rgb"
        Image.fromarray(rgb).save(scan_dir / "images" / f"{fid}.png")
        with open(scan_dir / "metadata" / "images" / f"{fid}.json", "w") as f:
            json.dump({"channel": "rgb", "camera": cam, "shot_id": f"{v:05d}"}, f)
        images.append({"id": fid, "file": f"{fid}.png"})
        for ch in CHANNELS:
            sid = f"{v:05d}_{ch}"
            Image.fromarray(vp_inputs["maps"][ch][v]).save(scan_dir / "Segmentation2D_1" / f"{sid}.png")
            with open(scan_dir / "metadata" / "Segmentation2D_1" / f"{sid}.json", "w") as f:
                json.dump({"channel": ch, "camera": cam, "shot_id": f"{v:05d}"}, f)
            seg.append({"id": sid, "file": f"{sid}.png"})
    with open(scan_dir / "metadata" / "images.json", "w") as f:
        json.dump({"bounding_box": vp_inputs["bbox"]}, f)
    with open(scan_dir / "files.json", "w") as f:
        json.dump({"filesets": [{"id": "images", "files": images}, {"id": "Segmentation2D_1", "files": seg}]}, f)

    # -- Masks (configs/geom_pipe_virtual.toml [Masks]: linear [0, 1, 0], dilation 1) ---------------------------
    masks_fs = tp2.run_masks(str(scan_dir), {"upstream_task": "ImagesFilesetExists", "type": "linear",
                                             "parameters": "[0,1,0]", "dilation": 1, "threshold": 0.3,
                                             "query": {"channel": "rgb"}})
    scan = fsdb_lite.Scan(str(scan_dir))
    mfiles = scan.get_fileset(masks_fs.id).get_files()
    assert len(mfiles) == n_views
    src = {f.id: f for f in scan.get_fileset("images").get_files()}
    for f in mfiles:
        want = mo.masks_filter(src[f.id].read_image(), "linear", [0, 1, 0], 0.3, mo.disk_sequence(1))
        assert np.array_equal(f.read_image(), want)
    # the reference's Masks task copies the camera along through romitask's FileByFileTask; do the same by hand
    for f in mfiles:
        f.set_metadata({"camera": src[f.id].get_metadata("camera"), "shot_id": src[f.id].get_metadata("shot_id")})

    # -- Voxels (carving from the masks just written) -----------------------------------------------------------
    vox = tcl.run_voxels(str(scan_dir), {"type": "carving", "voxel_size": vs, "camera_metadata": "camera",
                                         "upstream_mask": "Masks", "log": False})
    shape, origin = vp_shape_origin(vp_inputs["bbox"], vs)
    ora = BackprojectionOracle(shape, origin, vs, type="carving")
    want_vol = ora.process_fileset(fsdb_lite.Scan(str(scan_dir)).get_fileset(masks_fs.id).get_files(), "camera")
    got_vol = tcl.read_volume(vox.path())
    assert got_vol.shape == tuple(shape) and np.array_equal(got_vol, want_vol)
    assert (got_vol == 1).any() and (got_vol == -1).any()

    # -- PointCloud ---------------------------------------------------------------------------------------------
    pc = tp3.run_pointcloud(str(scan_dir), {"level_set_value": 1.0})
    pts, nrm, _ = tp3.read_ply(pc.path())
    want_p, want_n = vo.vol2pcd(want_vol, np.array(origin, dtype=np.float64), vs, 1.0)
    assert pts.shape == want_p.shape and len(pts) > 50
    np.testing.assert_allclose(pts, want_p, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(nrm, want_n, rtol=0, atol=1e-12)

    # -- SegmentedPointCloud ------------------------------------------------------------------------------------
    out = tp3.SegmentedPointCloud(str(scan_dir), use_colmap_poses=False).run()
    labels = sorted(c for c in CHANNELS if c != "background")
    segfiles = fsdb_lite.Scan(str(scan_dir)).get_fileset("Segmentation2D_1").get_files()
    views = [(f.get_metadata("channel"), f.get_metadata("camera"), f.read_image()) for f in segfiles]
    want_l, _ = vo.score_points(pts, views, labels)
    assert out.get_metadata("labels") == [labels[i] for i in want_l]
    assert len(set(out.get_metadata("labels"))) > 1

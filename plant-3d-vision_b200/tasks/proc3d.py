#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.tasks.proc3d
_______________________________

``PointCloud``: the consumer of the ``Voxels`` volume (reference: plant3dvision/tasks/proc3d.py:21-136), with the
same parameters and defaults, running ``plant3dvision_b200.proc3d`` (GPU) instead of scipy + joblib.  Only this
task of the reference module is mirrored: meshing, skeletons and the rest of tasks/proc3d.py are out of scope
(SURVEY §2).

luigi / romitask / plantdb / Open3D are not required: ``PointCloud(scan, **params).run()`` and
``run_pointcloud(scan_dir, config)`` work on a plain FSDB scan directory through ``fsdb_lite`` and write
``PointCloud.ply`` (binary little-endian, double x y z nx ny nz — the properties Open3D writes for a cloud with
normals; multiclass clouds also carry uchar red green blue) plus the file metadata the reference sets.
"""
import hashlib
import json
import logging
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from .. import fsdb_lite
from .. import proc3d
from .cl import _find_fileset, read_volume

logger = logging.getLogger(__name__)

#: parameter defaults of the reference task (tasks/proc3d.py:60-64)
DEFAULTS = dict(
    upstream_task="Voxels",
    level_set_value=1.0,
    background_prior=1.0,
    min_contrast=10.0,
    min_score=0.2,
)

#: label colours of romitask's PointCloudColorConfig defaults (used at tasks/proc3d.py:102, :118-122)
DEFAULT_COLORS = {"background": [0.0, 0.0, 0.0], "flower": [1.0, 0.0, 0.0], "peduncle": [0.0, 1.0, 0.0],
                  "pedicel": [0.0, 1.0, 0.0], "stem": [0.0, 0.0, 1.0], "leaf": [0.0, 1.0, 1.0],
                  "fruit": [1.0, 1.0, 0.0]}


def parse_params(section, strict=True):
    p = dict(DEFAULTS)
    for k, v in dict(section or {}).items():
        if k not in DEFAULTS:
            if strict:
                raise TypeError(f"Unknown PointCloud parameter '{k}'")
            logger.debug(f"Ignoring unknown [PointCloud] key '{k}'")
            continue
        p[k] = v
    for k in ("level_set_value", "background_prior", "min_contrast", "min_score"):
        p[k] = float(p[k])
    return p


def load_config(path):
    """Read a pipeline TOML and return its ``[PointCloud]`` parameters."""
    try:
        import tomllib
    except ImportError:  # pragma: no cover
        import tomli as tomllib
    with open(path, "rb") as f:
        cfg = tomllib.load(f)
    return parse_params(cfg.get("PointCloud", {}), strict=False)


def write_ply(path, points, normals, colors=None):
    """Binary little-endian PLY with the vertex properties Open3D's writer emits for such a cloud."""
    points = np.ascontiguousarray(points, dtype="<f8").reshape(-1, 3)
    normals = np.ascontiguousarray(normals, dtype="<f8").reshape(-1, 3)
    n = len(points)
    fields = [("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8")]
    header = ["ply", "format binary_little_endian 1.0", "comment Created by plant3dvision_b200", f"element vertex {n}"]
    header += [f"property double {name}" for name, _ in fields]
    if colors is not None:
        fields += [("red", "u1"), ("green", "u1"), ("blue", "u1")]
        header += ["property uchar red", "property uchar green", "property uchar blue"]
    header.append("end_header")
    rec = np.empty(n, dtype=np.dtype(fields))
    for a, name in enumerate(("x", "y", "z")):
        rec[name] = points[:, a]
        rec["n" + name] = normals[:, a]
    if colors is not None:
        c = np.clip(np.round(np.asarray(colors, dtype=np.float64).reshape(-1, 3) * 255.0), 0, 255).astype(np.uint8)
        rec["red"], rec["green"], rec["blue"] = c[:, 0], c[:, 1], c[:, 2]
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(rec.tobytes())


def read_ply(path):
    """Inverse of ``write_ply`` -> (points, normals, colors or None)."""
    with open(path, "rb") as f:
        names, n = [], 0
        while True:
            line = f.readline().decode("ascii").strip()
            if line.startswith("element vertex"):
                n = int(line.split()[-1])
            elif line.startswith("property"):
                _, t, name = line.split()
                names.append((name, "<f8" if t == "double" else "u1"))
            elif line == "end_header":
                break
        rec = np.frombuffer(f.read(), dtype=np.dtype(names), count=n)
    pts = np.stack([rec["x"], rec["y"], rec["z"]], axis=1)
    nrm = np.stack([rec["nx"], rec["ny"], rec["nz"]], axis=1)
    col = np.stack([rec["red"], rec["green"], rec["blue"]], axis=1) if "red" in rec.dtype.names else None
    return pts, nrm, col


class PointCloud:
    """Computes a point cloud from volumetric voxel data, single or multiclass (tasks/proc3d.py:21-136).

    ``scan`` is an ``fsdb_lite.Scan`` (or a path).  ``voxels_file`` may name the input file directly; otherwise the
    first file of the upstream ``Voxels*`` fileset is used, as ``RomiTask.input_file`` does."""

    def __init__(self, scan, voxels_file=None, colors=None, **params):
        self.scan = scan if hasattr(scan, "get_fileset") else fsdb_lite.Scan(scan)
        for k, v in parse_params(params).items():
            setattr(self, k, v)
        self._voxels_file = voxels_file
        self.colors = dict(DEFAULT_COLORS if colors is None else colors)

    @property
    def task_id(self):
        sig = json.dumps({k: getattr(self, k) for k in DEFAULTS}, sort_keys=True, default=str)
        return "PointCloud_" + hashlib.md5(sig.encode()).hexdigest()[:10]

    def input_file(self):
        if self._voxels_file is not None:
            return self._voxels_file
        fs = _find_fileset(self.scan, self.upstream_task)
        if fs is None or not fs.get_files():
            raise IOError(f"No output of upstream task '{self.upstream_task}' in scan '{self.scan.id}'")
        return fs.get_files()[0]

    @staticmethod
    def read_voxels(ifile):
        """tasks/proc3d.py:67-78: NPZ (one array per label) or a TIFF / NPY volume."""
        path = ifile.path() if callable(getattr(ifile, "path", None)) else ifile
        if str(path).endswith(".npz"):
            z = np.load(path)
            voxels = {k: z[k] for k in z.files}
            if len(voxels) == 1:
                return voxels[list(voxels.keys())[0]], False
            return voxels, True
        if str(path).endswith(".npy"):
            return np.load(path), False
        return read_volume(path), False

    def compute(self, device=None):
        """Everything of ``run`` except writing: -> (PointCloud, point_labels or None, colors or None, voxel_size)."""
        ifile = self.input_file()
        voxels, multiclass = self.read_voxels(ifile)
        origin = np.array(ifile.get_metadata("origin"))
        voxel_size = float(ifile.get_metadata("voxel_size"))
        if not multiclass:
            out = proc3d.vol2pcd(voxels, origin, voxel_size, self.level_set_value, device=device)
            return out, None, None, voxel_size
        pcd, point_labels = proc3d.pointcloud_from_voxels(voxels, origin, voxel_size, self.level_set_value,
                                                          self.background_prior, self.min_contrast, self.min_score,
                                                          device=device)
        rng = np.random.default_rng(0)  # the reference draws np.random.rand(3) for labels without a colour
        palette = {l: np.asarray(self.colors[l]) if l in self.colors else rng.random(3) for l in voxels.keys()}
        colors = np.array([palette[l] for l in point_labels], dtype=np.float64).reshape(-1, 3)
        return pcd, point_labels, colors, voxel_size

    def run(self, device=None):
        """Run and write ``PointCloud.ply`` + metadata (``{'labels': [...]}`` multiclass, else ``{'voxel_size'}``)."""
        pcd, point_labels, colors, voxel_size = self.compute(device=device)
        out_fs = self.scan.get_fileset(self.task_id, create=True)
        f = out_fs.create_file("PointCloud", "ply")
        write_ply(f.path(), pcd.points, pcd.normals, colors)
        if point_labels is not None:
            f.set_metadata({"labels": point_labels})
        else:
            f.set_metadata({"voxel_size": voxel_size})
        out_fs.set_metadata({"task_name": "PointCloud", "task_params": {k: getattr(self, k) for k in DEFAULTS}})
        return f


def run_pointcloud(scan_dir, config=None, **overrides):
    """``romi_run_task PointCloud <scan_dir> --config <toml>`` without luigi."""
    import os
    if isinstance(config, (str, os.PathLike)):
        params = load_config(config)
    else:
        params = parse_params(config, strict=False)
    params.update(overrides)
    return PointCloud(scan_dir, **params).run()


class SegmentedPointCloud:
    """Back-projects 2-D segmentation maps onto an existing point cloud (tasks/proc3d.py:139-253): every point is
    projected into every label map, the map values are summed per label and the point takes the arg-max label.
    The reference runs this as a Python loop over points inside a loop over files; here it is one kernel launch
    (``proc3d.score_points``).

    ``scan``: ``fsdb_lite.Scan`` or path; ``pcd_file`` / ``segmentation_fileset`` default to the first file of the
    ``PointCloud*`` fileset and to the ``Segmentation2D*`` fileset, as the reference's upstream tasks resolve.
    The reference builds its label list from a ``set`` (arbitrary order, which decides ties); here the labels are
    sorted, or given explicitly through ``labels``."""

    #: label-map pixels (bytes) decoded and uploaded per device call
    CHUNK_BYTES = 2 << 30

    def __init__(self, scan, pcd_file=None, segmentation_fileset=None, use_colmap_poses=True, labels=None,
                 colors=None, upstream_task="PointCloud", upstream_segmentation="Segmentation2D"):
        self.scan = scan if hasattr(scan, "get_fileset") else fsdb_lite.Scan(scan)
        self.use_colmap_poses = bool(use_colmap_poses)  # tasks/proc3d.py:160
        self.upstream_task = upstream_task
        self.upstream_segmentation = upstream_segmentation
        self._pcd_file = pcd_file
        self._seg_fs = segmentation_fileset
        self._labels = None if labels is None else list(labels)
        self.colors = dict(DEFAULT_COLORS if colors is None else colors)

    @property
    def task_id(self):
        sig = json.dumps({"use_colmap_poses": self.use_colmap_poses, "upstream_task": self.upstream_task,
                          "upstream_segmentation": self.upstream_segmentation}, sort_keys=True)
        return "SegmentedPointCloud_" + hashlib.md5(sig.encode()).hexdigest()[:10]

    def _inputs(self):
        fs = self._seg_fs if self._seg_fs is not None else _find_fileset(self.scan, self.upstream_segmentation)
        if fs is None:
            raise IOError(f"No output of '{self.upstream_segmentation}' in scan '{self.scan.id}'")
        pf = self._pcd_file
        if pf is None:
            pfs = _find_fileset(self.scan, self.upstream_task)
            if pfs is None or not pfs.get_files():
                raise IOError(f"No output of upstream task '{self.upstream_task}' in scan '{self.scan.id}'")
            pf = pfs.get_files()[0]
        return fs, pf

    def label_list(self, fs):
        """tasks/proc3d.py:192-201."""
        labels = set()
        for fi in fs.get_files():
            label = fi.get_metadata("channel")
            if label is not None:
                labels.add(label)
        labels = sorted(labels) if self._labels is None else list(self._labels)
        if self._labels is None:
            labels.remove("background")  # raises like the reference when there is no background channel
            if "rgb" in labels:
                labels.remove("rgb")
        return labels

    def compute(self, device=None):
        """-> (points, normals, pts_labels [N], labels list, scores [L, N])."""
        fs, pf = self._inputs()
        path = pf.path() if callable(getattr(pf, "path", None)) else pf
        pts, normals, _ = read_ply(path)
        labels = self.label_list(fs)
        key = "colmap_camera" if self.use_colmap_poses else "camera"  # tasks/proc3d.py:208-211
        todo = []
        for fi in fs.get_files():
            label, camera = fi.get_metadata("channel"), fi.get_metadata(key)
            if label not in labels:
                continue
            if camera is None:
                logger.warning("Could not get camera pose for view, skipping...")
                continue
            todo.append((label, camera, fi))
        # Label maps are decoded on a thread pool and scored in chunks of at most CHUNK_BYTES of pixels, so that host
        # and device memory stay bounded for large scans; the per-label sums are exact integers, so adding the
        # chunks' scores and taking the arg-max at the end equals one pass over all files.
        scores = np.zeros((len(labels), len(pts)))
        with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as pool:
            q = 0
            while labels and q < len(todo):
                first = np.asarray(todo[q][2].read_image())
                per = max(first.size, 1)
                e = min(len(todo), q + max(1, int(self.CHUNK_BYTES // per)))
                imgs = [first] + list(pool.map(lambda t: np.asarray(t[2].read_image()), todo[q + 1:e]))
                views = [(t[0], t[1], im) for t, im in zip(todo[q:e], imgs)]
                _, part, _ = proc3d.score_points(pts, views, labels, device=device)
                scores += part
                q = e
        pts_labels = np.argmax(scores, axis=0).flatten() if len(labels) else np.zeros(len(pts), dtype=np.int64)
        return pts, normals, pts_labels, labels, scores

    def run(self, device=None):
        """Writes ``SegmentedPointCloud.ply`` with per-label colours and the ``labels`` metadata
        (tasks/proc3d.py:231-253)."""
        pts, normals, pts_labels, labels, _ = self.compute(device=device)
        rng = np.random.default_rng(0)  # the reference draws np.random.rand(3) for labels without a colour
        palette = np.array([self.colors[l] if l in self.colors else rng.random(3) for l in labels],
                           dtype=np.float64).reshape(-1, 3)
        color_array = palette[pts_labels] if len(labels) else np.zeros((len(pts), 3))
        point_labels = [labels[i] for i in pts_labels.tolist()]
        out_fs = self.scan.get_fileset(self.task_id, create=True)
        f = out_fs.create_file("SegmentedPointCloud", "ply")
        write_ply(f.path(), pts, normals, color_array)
        f.set_metadata("labels", point_labels)
        out_fs.set_metadata({"task_name": "SegmentedPointCloud",
                             "task_params": {"use_colmap_poses": self.use_colmap_poses}})
        return f

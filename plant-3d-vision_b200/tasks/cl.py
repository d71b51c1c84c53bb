#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.tasks.cl
___________________________

``Voxels``: same parameters, defaults, bounding-box / shape / label resolution and outputs as the
reference task (plant3dvision/tasks/cl.py:18-186), running on ``plant3dvision_b200.cl.Backprojection``.

luigi / romitask / plantdb are not required: ``Voxels(scan, **params).run()`` and
``run_voxels(scan_dir, config)`` work on a plain FSDB scan directory through ``fsdb_lite``.  The
``[Voxels]`` section of the reference's TOML configs is accepted verbatim, including list / dict
parameters given as JSON strings (e.g. ``labels = "[\\"background\\"]"``, configs/ml_pipe_real.toml:46)
and the nested ``[Voxels.bounding_box]`` table (configs/geom_pipe_real.toml:36-39).
"""
import hashlib
import json
import logging
import os
import sys

import numpy as np

from .. import fsdb_lite
from ..npz_parallel import savez_compressed_parallel

logger = logging.getLogger(__name__)

#: parameter defaults of the reference task (tasks/cl.py:79-91)
DEFAULTS = dict(
    upstream_mask="Masks",
    upstream_colmap="Colmap",
    query={},
    camera_metadata="colmap_camera",
    voxel_size=1.0,
    type="carving",
    log=True,
    invert=False,
    labels=[],
    bounding_box=None,
)


def parse_params(section, strict=True):
    """``[Voxels]`` TOML section -> parameter dict with the reference's defaults.  luigi's config
    ingestion hands list / dict parameters over as JSON strings; both forms are accepted."""
    p = dict(DEFAULTS)
    for k, v in dict(section or {}).items():
        if k not in DEFAULTS and not strict:
            # luigi's config ingestion ignores keys that are not parameters of the task
            # (e.g. `use_colmap_poses` in configs/ml_pipe_real.toml:41)
            logger.debug(f"Ignoring unknown [Voxels] key '{k}'")
            continue
        if k in ("query", "labels", "bounding_box") and isinstance(v, str):
            v = json.loads(v)
        p[k] = v
    p["voxel_size"] = float(p["voxel_size"])
    p["log"] = bool(p["log"])
    p["invert"] = bool(p["invert"])
    p["labels"] = list(p["labels"] or [])
    p["query"] = dict(p["query"] or {})
    if p["bounding_box"] is not None:
        p["bounding_box"] = {a: list(p["bounding_box"][a]) for a in "xyz"}
    return p


def load_config(path):
    """Read a pipeline TOML and return its ``[Voxels]`` parameters."""
    try:
        import tomllib
    except ImportError:  # pragma: no cover
        import tomli as tomllib
    with open(path, "rb") as f:
        cfg = tomllib.load(f)
    return parse_params(cfg.get("Voxels", {}), strict=False)


def grid_from_bbox(bounding_box, voxel_size, displacement=None):
    """Shape and origin of the voxel array (tasks/cl.py:127-147), in python floats like the task."""
    x_min, x_max = bounding_box["x"]
    y_min, y_max = bounding_box["y"]
    z_min, z_max = bounding_box["z"]
    if displacement is not None:
        x_min += displacement["dx"]
        x_max += displacement["dx"]
        y_min += displacement["dy"]
        y_max += displacement["dy"]
        z_min += displacement["dz"]
        z_max += displacement["dz"]
    nx = int((x_max - x_min) / voxel_size) + 1
    ny = int((y_max - y_min) / voxel_size) + 1
    nz = int((z_max - z_min) / voxel_size) + 1
    return [nx, ny, nz], [x_min, y_min, z_min]


def _find_fileset(scan, task_name):
    """The fileset an upstream task wrote: 'images' for ImagesFilesetExists, else the first
    fileset named after the task (romitask names them ``<Task>_<params hash>``)."""
    if task_name in ("ImagesFilesetExists", "images"):
        return scan.get_fileset("images")
    for fs in scan.get_filesets():
        if fs.id == task_name or fs.id.startswith(task_name + "_"):
            return fs
    return None


class Voxels:
    """Computes a volume from backprojection of 2D segmented images.

    Parameters are those of the reference task; ``scan`` is an ``fsdb_lite.Scan`` (or a path)."""

    def __init__(self, scan, masks_fileset=None, colmap_fileset=None, **params):
        self.scan = scan if hasattr(scan, "get_fileset") else fsdb_lite.Scan(scan)
        self._colmap_fileset = colmap_fileset
        unknown = set(params) - set(DEFAULTS)
        if unknown:
            raise TypeError(f"Unknown Voxels parameters: {sorted(unknown)}")
        for k, v in parse_params(params).items():
            setattr(self, k, v)
        self._masks_fileset = masks_fileset

    @property
    def task_id(self):
        sig = json.dumps({k: getattr(self, k) for k in DEFAULTS}, sort_keys=True, default=str)
        return "Voxels_" + hashlib.md5(sig.encode()).hexdigest()[:10]

    def resolve_bounding_box(self, masks_fileset):
        """Precedence of tasks/cl.py:107-118: parameter, scan metadata, Colmap fileset, images fileset."""
        bbox = self.bounding_box
        if bbox is None:
            bbox = self.scan.get_metadata("bounding_box")
        if bbox is None and self.upstream_colmap == "Colmap":
            colmap_fs = self._colmap_fileset or _find_fileset(self.scan, "Colmap")
            if colmap_fs is not None:
                bbox = colmap_fs.get_metadata("bounding_box")
        if bbox is None:
            images = self.scan.get_fileset("images")
            if images is not None:
                bbox = images.get_metadata("bounding_box")
        return bbox

    def compute(self, backprojection_class=None, **bp_kwargs):
        """Everything of ``run`` up to (not including) writing: returns (vol, labels, origin).
        ``backprojection_class``: the engine, ``plant3dvision_b200.cl.Backprojection`` by default;
        ``plant3dvision_b200.parallel.DistributedBackprojection`` under ``torchrun`` shards the grid or the views
        over the ranks (its keyword arguments - ``sharding=...`` - go through ``bp_kwargs``)."""
        if backprojection_class is None:
            from ..cl import Backprojection as backprojection_class
        Backprojection = backprojection_class
        masks_fileset = self._masks_fileset or _find_fileset(self.scan, self.upstream_mask)
        if masks_fileset is None:
            raise IOError(f"No fileset found for upstream task '{self.upstream_mask}' in scan '{self.scan.id}'")
        masks_files = masks_fileset.get_files(query=self.query)
        logger.info(f"Processing a list of {len(masks_files)} mask files...")

        bounding_box = self.resolve_bounding_box(masks_fileset)
        if bounding_box is None:
            logger.critical(f"Could not obtain valid bounding-box for {self.scan.id}!")
            sys.exit("Error with bounding-box definition!")
        logger.info(f"Bounding-box to use: {bounding_box}")

        displacement = self.scan.get_metadata("displacement")
        if displacement is None:
            logger.warning("No 'displacement' found in scan metadata!")
        shape, origin = grid_from_bbox(bounding_box, self.voxel_size, displacement)

        if len(self.labels) == 0:
            labels = masks_fileset.get_metadata("label_names", default=None)
            if labels is None or len(labels) == 0:
                logger.warning("No metadata 'label_names' in `masks_fileset`!")
                labels = None
        else:
            labels = list(self.labels)

        sc = Backprojection(shape=shape, origin=origin, voxel_size=float(self.voxel_size), type=str(self.type),
                            labels=labels, log=bool(self.log), **bp_kwargs)
        # The task tail (tasks/cl.py:166-174) is fused into the read-back: min / max of the raw volume replace the
        # np.unique sort of :168, np.exp and the clip to 1 of :172-174 run on the device.
        stats = {}
        vol = sc.process_fileset(masks_files, str(self.camera_metadata), bool(self.invert),
                                 exp_clip=bool(self.log and self.type == "averaging"), stats=stats)
        if labels is None:
            vol = vol.copy()  # detach from the pinned buffer before the device handle goes away
        self._rank = int(getattr(sc, "rank", 0))
        sc.close()
        logger.debug(f"Voxel volume shape: {vol.shape}")
        if vol.size and stats["min"] == stats["max"]:
            logger.warning("There is something WRONG with the volume!")
        return vol, labels, origin

    def run(self, backprojection_class=None, **bp_kwargs):
        """Run and write the output fileset: ``Voxels.npz`` (one array per label) or ``Voxels.tiff``
        / ``Voxels.npy`` plus the file metadata ``{voxel_size, origin}`` (tasks/cl.py:176-186).  With a distributed
        engine every rank computes and rank 0 writes (the others return ``None``)."""
        vol, labels, origin = self.compute(backprojection_class, **bp_kwargs)
        if getattr(self, "_rank", 0) != 0:
            return None
        out_fs = self.scan.get_fileset(self.task_id, create=True)
        if labels is not None:
            f = out_fs.create_file("Voxels", "npz")
            savez_compressed_parallel(f.path(), {label: vol[i, :] for i, label in enumerate(labels)})
        else:
            f = out_fs.create_file("Voxels", "tiff")
            write_volume(f.path(), vol)
        f.set_metadata({'voxel_size': self.voxel_size, 'origin': [float(o) for o in origin]})
        out_fs.set_metadata({"task_name": "Voxels", "task_params": {k: getattr(self, k) for k in DEFAULTS}})
        return f


#: volumes at least this large are written as BigTIFF by the PIL fallback of ``write_volume``
BIGTIFF_BYTES = 2 ** 32 - 2 ** 24


def write_volume(path, vol):
    """Multi-page TIFF, one page per axis-0 plane - the layout imageio gives plantdb's
    ``io.write_volume`` (io.py:491-525).  Uses tifffile / imageio when present, else PIL."""
    try:
        import tifffile
        tifffile.imwrite(path, vol)
        return
    except ImportError:
        pass
    from PIL import Image
    pages = [Image.fromarray(np.ascontiguousarray(vol[i])) for i in range(vol.shape[0])]
    # classic TIFF holds 32-bit offsets: a 1024^3 int32 volume (4.29 GB) needs BigTIFF, which is what
    # tifffile / imageio switch to on their own
    big = vol.nbytes + 4096 * len(pages) >= BIGTIFF_BYTES
    pages[0].save(path, format="TIFF", save_all=True, append_images=pages[1:], **({"big_tiff": True} if big else {}))


def read_volume(path):
    """Inverse of ``write_volume`` (plantdb io.read_volume)."""
    try:
        import tifffile
        return tifffile.imread(path)
    except ImportError:
        pass
    from PIL import Image, ImageSequence
    with Image.open(path) as im:
        return np.stack([np.asarray(p) for p in ImageSequence.Iterator(im)])


def run_voxels(scan_dir, config=None, backprojection_class=None, bp_kwargs=None, **overrides):
    """``romi_run_task Voxels <scan_dir> --config <toml>`` without luigi: ``config`` is a TOML path
    or an already parsed ``[Voxels]`` dict.  ``backprojection_class`` / ``bp_kwargs``: see ``Voxels.compute``
    (e.g. ``DistributedBackprojection`` with ``{"sharding": "stripes"}`` under ``torchrun``)."""
    if isinstance(config, (str, os.PathLike)):
        params = load_config(config)
    else:
        params = parse_params(config, strict=False)
    params.update(overrides)
    return VoxelsRunner(scan_dir, **params).run(backprojection_class, **(bp_kwargs or {}))


# ---------------------------------------------------------------------------------------------
# luigi / romitask registration
# ---------------------------------------------------------------------------------------------
#: the plain, dependency-free runner under a name that never changes
VoxelsRunner = Voxels


def make_luigi_voxels():
    """The ``Voxels`` luigi task of the reference (plant3dvision/tasks/cl.py:18-186: same task family, same
    ``luigi.*Parameter`` attributes and defaults, same ``requires()``), running on this package's ``Backprojection``.
    Needs luigi + romitask (+ plantdb for the writers); ``romitask/modules.py:43`` maps the name ``Voxels`` to a module
    path - pointing it at ``plant3dvision_b200.tasks.cl`` is the whole integration."""
    import luigi
    from romitask import RomiTask

    def _upstream(name, module):
        try:
            mod = __import__(module, fromlist=[name])
            return luigi.TaskParameter(default=getattr(mod, name))
        except ImportError:  # plant3dvision itself not installed: the task must then be given explicitly
            return luigi.TaskParameter()

    class Voxels(RomiTask):
        upstream_task = None  # override default attribute from ``RomiTask``
        upstream_mask = _upstream("Masks", "plant3dvision.tasks.proc2d")
        upstream_colmap = _upstream("Colmap", "plant3dvision.tasks.colmap")

        query = luigi.DictParameter(default={})
        camera_metadata = luigi.Parameter(default='colmap_camera')
        voxel_size = luigi.FloatParameter(default=1.0)
        type = luigi.Parameter(default="carving")
        log = luigi.BoolParameter(default=True)

        invert = luigi.BoolParameter(default=False)
        labels = luigi.ListParameter(default=[])
        bounding_box = luigi.DictParameter(default=None)

        def requires(self):
            if self.upstream_colmap.get_task_family() == 'Colmap':
                return {'masks': self.upstream_mask(), 'colmap': self.upstream_colmap()}
            else:
                return {'masks': self.upstream_mask()}

        def run(self):
            masks_fileset = self.input()['masks'].get()
            uses_colmap = self.upstream_colmap.get_task_family() == 'Colmap'
            colmap_fileset = self.input()['colmap'].get() if uses_colmap else None
            runner = VoxelsRunner(
                masks_fileset.scan, masks_fileset=masks_fileset, colmap_fileset=colmap_fileset,
                upstream_colmap="Colmap" if uses_colmap else self.upstream_colmap.get_task_family(),
                query=dict(self.query), camera_metadata=str(self.camera_metadata), voxel_size=float(self.voxel_size),
                type=str(self.type), log=bool(self.log), invert=bool(self.invert), labels=list(self.labels),
                bounding_box=None if self.bounding_box is None else {k: list(v) for k, v in dict(self.bounding_box).items()})
            vol, labels, origin = runner.compute()
            outfile = self.output_file()
            from plantdb import io
            if labels is not None:
                io.write_npz(outfile, {label: vol[i, :] for i, label in enumerate(labels)})
            else:
                io.write_volume(outfile, vol)
            outfile.set_metadata({'voxel_size': self.voxel_size, 'origin': [float(o) for o in origin]})

    return Voxels


try:  # registered only where the reference's task framework exists; everywhere else `Voxels` is the plain runner
    import luigi as _luigi  # noqa: F401
    import romitask as _romitask  # noqa: F401
except ImportError:
    pass
else:
    Voxels = make_luigi_voxels()

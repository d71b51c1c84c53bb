#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.tasks.proc2d
_______________________________

``Masks``: the producer of the ``Voxels`` task's input (reference: plant3dvision/tasks/proc2d.py:140-248), same
parameters and defaults, same output files and metadata, with the per-image host work (skimage rescale, numpy index,
threshold, skimage dilation) replaced by one batched GPU call (``plant3dvision_b200.proc2d.masks``) and the PNG
decode / encode spread over a thread pool.  Only this task of the reference module is mirrored (``Undistorted`` needs
OpenCV, ``Segmentation2D`` a trained network: out of scope).
"""
import hashlib
import json
import logging
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from .. import fsdb_lite
from .. import proc2d
from .cl import _find_fileset

logger = logging.getLogger(__name__)

#: parameter defaults of the reference task (tasks/proc2d.py:207-211; query from FileByFileTask)
DEFAULTS = dict(
    upstream_task="Undistorted",
    query={},
    type="linear",
    parameters=[0, 1, 0],
    threshold=0.3,
    dilation=0,
)

#: images per GPU call (bounds device memory: 64 x 24 Mpx x 3 B = 4.6 GB)
BATCH = 64


def parse_params(section, strict=True):
    p = dict(DEFAULTS)
    for k, v in dict(section or {}).items():
        if k not in DEFAULTS:
            if strict:
                raise TypeError(f"Unknown Masks parameter '{k}'")
            logger.debug(f"Ignoring unknown [Masks] key '{k}'")
            continue
        p[k] = v
    if isinstance(p["parameters"], str):  # the reference's TOMLs spell the list as a string: parameters = "[0,1,0]"
        p["parameters"] = json.loads(p["parameters"])
    if isinstance(p["query"], str):
        p["query"] = json.loads(p["query"].replace("'", '"'))
    p["parameters"] = list(p["parameters"])
    p["threshold"] = float(p["threshold"])
    p["dilation"] = int(p["dilation"])
    return p


def load_config(path):
    """Read a pipeline TOML and return its ``[Masks]`` parameters."""
    try:
        import tomllib
    except ImportError:  # pragma: no cover
        import tomli as tomllib
    with open(path, "rb") as f:
        cfg = tomllib.load(f)
    return parse_params(cfg.get("Masks", {}), strict=False)


def _read(fi):
    from ..cl import _read_image
    return np.asarray(_read_image(fi))


def _write_png(path, mask):
    from PIL import Image
    Image.fromarray(mask).save(path)


class Masks:
    """Compute binary masks from RGB images (tasks/proc2d.py:140-248)."""

    def __init__(self, scan, images_fileset=None, dilation_sequence=None, **params):
        self.scan = scan if hasattr(scan, "get_fileset") else fsdb_lite.Scan(scan)
        for k, v in parse_params(params).items():
            setattr(self, k, v)
        if self.type not in ("linear", "excess_green"):
            raise Exception(f"Unknown masking type '{self.type}'!")  # tasks/proc2d.py:221-222
        self._images_fileset = images_fileset
        self._sequence = dilation_sequence  # explicit footprints instead of skimage's disk(n, 'sequence')

    @property
    def task_id(self):
        sig = json.dumps({k: getattr(self, k) for k in DEFAULTS}, sort_keys=True, default=str)
        return "Masks_" + hashlib.md5(sig.encode()).hexdigest()[:10]

    def input_files(self):
        fs = self._images_fileset
        if fs is None:
            fs = _find_fileset(self.scan, self.upstream_task) or _find_fileset(self.scan, "images")
        if fs is None:
            raise IOError(f"No fileset found for upstream task '{self.upstream_task}' in scan '{self.scan.id}'")
        return fs.get_files(query=self.query)

    def compute(self, device=None, threads=None):
        """-> list of (input File, uint8 mask) in file order."""
        files = self.input_files()
        dil = self._sequence if self._sequence is not None else self.dilation
        out = []
        with ThreadPoolExecutor(max_workers=threads or min(16, os.cpu_count() or 1)) as pool:
            for o in range(0, len(files), BATCH):
                chunk = files[o:o + BATCH]
                imgs = list(pool.map(_read, chunk))
                # images of one scan share a size; group by shape anyway so that a stray file cannot break the batch
                by_shape = {}
                for i, im in enumerate(imgs):
                    by_shape.setdefault(im.shape, []).append(i)
                res = [None] * len(chunk)
                for idx in by_shape.values():
                    ms = proc2d.masks([imgs[i] for i in idx], self.type, self.parameters, self.threshold, dil,
                                      device=device)
                    for i, m in zip(idx, ms):
                        res[i] = m
                out += list(zip(chunk, res))
        return out

    def file_metadata(self):
        """tasks/proc2d.py:242-247."""
        md = {"upstream_task": str(self.upstream_task), "filter": str(self.type), "threshold": self.threshold,
              "dilation": self.dilation}
        if self.type == "linear":
            md.update({"linear_coeff": list(self.parameters)})
        if self.query != {}:
            md.update({"query": dict(self.query)})
        return {"Masks": md}

    def run(self, device=None, threads=None):
        """Write one PNG mask per input image under the same file id, with the reference's metadata."""
        results = self.compute(device=device, threads=threads)
        out_fs = self.scan.get_fileset(self.task_id, create=True)
        md = self.file_metadata()
        outs = []
        with ThreadPoolExecutor(max_workers=threads or min(16, os.cpu_count() or 1)) as pool:
            jobs = []
            for fi, mask in results:
                f = out_fs.create_file(fi.id, "png")
                jobs.append(pool.submit(_write_png, f.path(), mask))
                outs.append(f)
            for j in jobs:
                j.result()
        for f in outs:
            f.set_metadata(dict(md))
        out_fs.set_metadata({"task_name": "Masks", "task_params": {k: getattr(self, k) for k in DEFAULTS}})
        return out_fs


def run_masks(scan_dir, config=None, **overrides):
    """``romi_run_task Masks <scan_dir> --config <toml>`` without luigi."""
    if isinstance(config, (str, os.PathLike)):
        params = load_config(config)
    else:
        params = parse_params(config, strict=False)
    params.update(overrides)
    return Masks(scan_dir, **params).run()

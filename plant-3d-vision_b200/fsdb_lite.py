"""Minimal reader/writer of the on-disk layout of a plantdb FSDB *scan directory*, enough for the
``Voxels`` task to run without plantdb installed (reference layout: plantdb/src/plantdb/fsdb.py:
``<scan>/files.json``, ``<scan>/metadata/metadata.json`` (scan), ``<scan>/metadata/<fileset>.json``
and ``<scan>/metadata/<fileset>/<file>.json``; query filter semantics of fsdb.py:2735-2791).

Objects are duck-typed like plantdb's ``Scan`` / ``Fileset`` / ``File``: ``id``,
``get_metadata(key=None, default=None)``, ``get_files(query=None)``, ``path()``; ``Backprojection``
accepts either these or real plantdb objects.
"""
import json
import os

import numpy as np


def _load_json(path):
    if os.path.isfile(path):
        with open(path) as f:
            return json.load(f)
    return {}


def _dump_json(path, obj):
    """Write-then-rename, so that a concurrent reader (another rank of a distributed task reading the scan while
    rank 0 writes its output fileset) sees the old or the new file, never a truncated one."""
    tmp = f"{path}.tmp{os.getpid()}"
    with open(tmp, "w") as f:
        json.dump(obj, f, indent=4)
    os.replace(tmp, path)


def _get(md, key, default):
    if key is None:
        return md
    return md.get(key, default)


class File:
    def __init__(self, fileset, fid, filename):
        self.fileset = fileset
        self.id = fid
        self.filename = filename
        self._md = None

    @property
    def metadata(self):
        if self._md is None:
            self._md = _load_json(os.path.join(self.fileset.scan.dir, "metadata", self.fileset.id, self.id + ".json"))
        return self._md

    def get_metadata(self, key=None, default=None):
        return _get(self.metadata, key, default)

    def set_metadata(self, data, value=None):
        if not isinstance(data, dict):
            data = {data: value}
        self.metadata.update(data)
        d = os.path.join(self.fileset.scan.dir, "metadata", self.fileset.id)
        os.makedirs(d, exist_ok=True)
        _dump_json(os.path.join(d, self.id + ".json"), self._md)

    def path(self):
        return os.path.join(self.fileset.scan.dir, self.fileset.id, self.filename)

    def read_image(self):
        # plantdb.io.read_image (io.py:343-378) returns imageio's array: 2-D for 'L' PNGs
        from PIL import Image
        with Image.open(self.path()) as im:
            return np.asarray(im)


class Fileset:
    def __init__(self, scan, fsid, file_infos):
        self.scan = scan
        self.id = fsid
        self.files = [File(self, fi["id"], fi["file"]) for fi in file_infos]
        self._md = None

    @property
    def metadata(self):
        if self._md is None:
            self._md = _load_json(os.path.join(self.scan.dir, "metadata", self.id + ".json"))
        return self._md

    def get_metadata(self, key=None, default=None):
        return _get(self.metadata, key, default)

    def set_metadata(self, data, value=None):
        if not isinstance(data, dict):
            data = {data: value}
        self.metadata.update(data)
        os.makedirs(os.path.join(self.scan.dir, "metadata"), exist_ok=True)
        _dump_json(os.path.join(self.scan.dir, "metadata", self.id + ".json"), self._md)

    def get_files(self, query=None):
        """All files whose metadata match every key/value of ``query`` (fsdb.py:2735-2791)."""
        if not query:
            return list(self.files)
        return [f for f in self.files if all(f.get_metadata(k) == v for k, v in dict(query).items())]

    def get_file(self, fid):
        for f in self.files:
            if f.id == fid:
                return f
        return None

    def create_file(self, fid, ext):
        """Register ``<fileset>/<fid>.<ext>`` in files.json and return the File."""
        f = self.get_file(fid)
        if f is None:
            f = File(self, fid, f"{fid}.{ext}")
            self.files.append(f)
        else:
            f.filename = f"{fid}.{ext}"
        os.makedirs(os.path.join(self.scan.dir, self.id), exist_ok=True)
        self.scan._save()
        return f


class Scan:
    def __init__(self, scan_dir):
        self.dir = os.path.abspath(str(scan_dir))
        self.id = os.path.basename(self.dir.rstrip("/"))
        info = _load_json(os.path.join(self.dir, "files.json"))
        self.filesets = [Fileset(self, fs["id"], fs.get("files", [])) for fs in info.get("filesets", [])]
        self._md = None

    @property
    def metadata(self):
        if self._md is None:
            self._md = _load_json(os.path.join(self.dir, "metadata", "metadata.json"))
        return self._md

    def get_metadata(self, key=None, default=None):
        return _get(self.metadata, key, default)

    def get_fileset(self, fsid, create=False):
        for fs in self.filesets:
            if fs.id == fsid:
                return fs
        if create:
            fs = Fileset(self, fsid, [])
            self.filesets.append(fs)
            os.makedirs(os.path.join(self.dir, fsid), exist_ok=True)
            self._save()
            return fs
        return None

    def get_filesets(self):
        return list(self.filesets)

    def _save(self):
        info = {"filesets": [{"id": fs.id, "files": [{"id": f.id, "file": f.filename} for f in fs.files]}
                             for fs in self.filesets]}
        _dump_json(os.path.join(self.dir, "files.json"), info)

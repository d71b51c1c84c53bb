"""TEST INFRASTRUCTURE — not product code.

CPU oracle for the voxel back-projection path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this module; the product package never does.

It restates the *host* side of the reference's ``Backprojection``
(/root/reference/plant3dvision/cl.py:118-311) in numpy and runs the per-view
arithmetic in one of three compiled back ends:

``restatement``  ``oracle/libbp_oracle.so``  - our C restatement (``bp_oracle.c``)
``ref_off``      ``oracle/_ref/libbp_ref_off.so`` - the reference's own
                 ``kernels/backprojection.c`` compiled unmodified behind
                 ``ref_shim.h`` with ``-ffp-contract=off`` (canonical oracle)
``ref_fma``      same, FMA-contracted (reported next to results, never gated on)

PARITY PINNING: the reference's tests hold no numeric vector for this path
(tests/unit/test_cl.py:4-9 constructs only), and pyopencl cannot be imported here.
The restatement is pinned against ``ref_off`` (the reference's own kernel source) on
the reference's ``virtual_plant`` fixture; see ``tests/golden/make_golden.py``.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
EPS = 1e-10  # cl.py:35

_F = ctypes.POINTER(ctypes.c_float)
_I = ctypes.POINTER(ctypes.c_int32)
_B = ctypes.POINTER(ctypes.c_uint8)
_libs = {}


def build(quiet=True):
    """Compile the restatement (and oracle/_ref when /root/reference is present)."""
    subprocess.run(["make", "-C", _HERE], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def have_backend(name):
    return os.path.exists(_lib_path(name))


def _lib_path(name):
    if name == "restatement":
        return os.path.join(_HERE, "libbp_oracle.so")
    if name in ("ref_off", "ref_fma"):
        return os.path.join(_HERE, "_ref", "libbp_%s.so" % name)
    raise ValueError(name)


def _lib(name):
    if name not in _libs:
        path = _lib_path(name)
        if not os.path.exists(path):
            if name == "restatement":
                build()
            else:
                raise FileNotFoundError(path)
        _libs[name] = ctypes.CDLL(path)
    return _libs[name]


def _p(a, t):
    return a.ctypes.data_as(t)


def img_as_float32(mask):
    """What ``skimage.util.img_as_float32`` returns for the dtypes masks come in
    (cl.py:205-206): uint8 -> ``np.multiply(m, 1./255, dtype=float32)``; uint16 likewise
    with 65535; bool -> 0./1.; float -> float32 cast."""
    mask = np.asarray(mask)
    if mask.dtype == np.float32:
        return mask
    if mask.dtype == np.bool_:
        return mask.astype(np.float32)
    if mask.dtype.kind == "u":
        return np.multiply(mask, 1.0 / np.iinfo(mask.dtype).max, dtype=np.float32)
    if mask.dtype.kind == "f":
        return mask.astype(np.float32)
    raise TypeError("oracle: unsupported mask dtype %s" % mask.dtype)


class BackprojectionOracle:
    """Same constructor / methods as the reference class (cl.py:118-311)."""

    def __init__(self, shape, origin, voxel_size, type="carving", default_value=0, labels=None,
                 log=False, backend="restatement", i_offset=0):
        self.shape = [int(s) for s in shape]
        self.origin = origin
        self.voxel_size = voxel_size
        self.default_value = default_value
        self.log = log
        self.labels = labels
        self.backend = backend
        self.i_offset = int(i_offset)
        if type == "carving":
            self.dtype = np.int32
        elif type == "averaging":
            self.dtype = np.float32
        else:
            raise ValueError(f"Unknown kernel type {type}, valid values are 'averaging' or 'carving'!")
        self.lib = _lib(backend)
        if backend != "restatement" and self.i_offset:
            raise ValueError("i_offset is a restatement-only extension")
        self.values_h = None
        self.init_buffers()

    def init_buffers(self):
        self.values_h = self.default_value * np.ones(self.shape, dtype=self.dtype)  # cl.py:173
        # cl.py:181-187: volinfo = [*origin, voxel_size] as float32; shape as int32
        self.volinfo = np.array([*self.origin, self.voxel_size], dtype=np.float32)
        self.shape_i = np.array(self.shape, dtype=np.int32)

    def process_view(self, intrinsics, rot, tvec, mask):
        mask = np.asarray(mask)
        if self.dtype == np.float32 and mask.dtype != np.float32:  # cl.py:205-206
            mask = img_as_float32(mask)
        if self.log and self.dtype == np.float32:  # cl.py:207-208
            mask = np.log(EPS + mask)
        intr = np.ascontiguousarray(intrinsics, dtype=np.float32)
        rot = np.ascontiguousarray(rot, dtype=np.float32).reshape(-1)
        tvec = np.ascontiguousarray(tvec, dtype=np.float32)
        H, W = mask.shape[:2]
        if self.backend == "restatement":
            origin = np.ascontiguousarray(self.volinfo[:3])
            vs = ctypes.c_float(float(self.volinfo[3]))
            if self.dtype == np.int32:
                # cl.py:215 casts to int32 first; the kernel then only tests `== 0`
                m = np.ascontiguousarray(np.ascontiguousarray(mask, dtype=np.int32) != 0, dtype=np.uint8)
                self.lib.oracle_carve_view(_p(m, _B), H, W, _p(self.values_h, _I), _p(intr, _F),
                                           _p(rot, _F), _p(tvec, _F), _p(origin, _F), vs,
                                           _p(self.shape_i, _I), self.i_offset)
            else:
                m = np.ascontiguousarray(mask, dtype=np.float32)
                self.lib.oracle_average_view(_p(m, _F), H, W, _p(self.values_h, _F), _p(intr, _F),
                                             _p(rot, _F), _p(tvec, _F), _p(origin, _F), vs,
                                             _p(self.shape_i, _I), self.i_offset)
        else:
            mask_h = np.ascontiguousarray(mask, dtype=self.dtype)  # cl.py:215
            if self.dtype == np.int32:
                self.lib.ref_carve_view(_p(mask_h, _I), H, W, _p(self.values_h, _I), _p(intr, _F),
                                        _p(rot, _F), _p(tvec, _F), _p(self.volinfo, _F),
                                        _p(self.shape_i, _I))
            else:
                self.lib.ref_average_view(_p(mask_h, _F), H, W, _p(self.values_h, _F), _p(intr, _F),
                                          _p(rot, _F), _p(tvec, _F), _p(self.volinfo, _F),
                                          _p(self.shape_i, _I))

    def process_view_u8_lut(self, intrinsics, rot, tvec, mask_u8):
        """Averaging on a uint8 map through the 256-entry table of what
        cl.py:205-208 gives per byte value (restatement only; large workloads)."""
        assert self.backend == "restatement" and self.dtype == np.float32
        lut = img_as_float32(np.arange(256, dtype=np.uint8))
        if self.log:
            lut = np.log(EPS + lut)
        lut = np.ascontiguousarray(lut, dtype=np.float32)
        intr = np.ascontiguousarray(intrinsics, dtype=np.float32)
        rot = np.ascontiguousarray(rot, dtype=np.float32).reshape(-1)
        tvec = np.ascontiguousarray(tvec, dtype=np.float32)
        m = np.ascontiguousarray(mask_u8, dtype=np.uint8)
        H, W = m.shape
        origin = np.ascontiguousarray(self.volinfo[:3])
        self.lib.oracle_average_view_lut(_p(m, _B), _p(lut, _F), H, W, _p(self.values_h, _F),
                                         _p(intr, _F), _p(rot, _F), _p(tvec, _F), _p(origin, _F),
                                         ctypes.c_float(float(self.volinfo[3])),
                                         _p(self.shape_i, _I), self.i_offset)

    def get_values(self):
        return self.values_h.reshape(self.shape)  # cl.py:229-232

    def clear(self):
        self.values_h = self.default_value * np.ones(self.shape).astype(self.dtype)  # cl.py:309

    def process_fileset(self, fs, camera_metadata, invert=False):  # cl.py:234-257
        if self.labels is not None:
            result = np.zeros((len(self.labels), *self.shape))
            for i, label in enumerate(self.labels):
                if i != 0:
                    self.clear()
                result[i, :] = self.process_label(fs, camera_metadata, label, invert)
            return result
        return self.process_label(fs, camera_metadata, None, invert=invert)

    def process_label(self, fs, camera_metadata, label=None, invert=False):  # cl.py:259-305
        if hasattr(fs, "get_files"):
            fs = fs.get_files()
        for fi in fs:
            if label is not None and fi.get_metadata("channel") != label:
                continue
            cam = fi.get_metadata(camera_metadata, default=None)
            if cam is None:
                continue
            intrinsics = np.array(cam["camera_model"]["params"][0:4], dtype=np.float32)
            rot = np.array(sum(cam["rotmat"], []), dtype=np.float32)
            tvec = np.array(cam["tvec"], dtype=np.float32)
            mask = fi.read_image()
            if invert:
                mask = np.invert(mask)
            self.process_view(intrinsics, rot, tvec, mask)
        return self.get_values()

#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.cl
_____________________

Drop-in for the ``Backprojection`` class of ``plant3dvision.cl`` (reference:
plant3dvision/cl.py:47-311): same constructor, attributes, methods, return dtypes and
shapes.  The OpenCL context / programs of the reference (cl.py:29-44) are replaced by
``libbp_b200.so`` (sm_100a CUDA behind the C ABI of ``include/bp_b200.h``).

Differences a caller can observe, all additive:

* ``process_label`` / ``process_fileset`` hand all views of a label to ONE fused kernel
  (``bp_process_views``) instead of one launch + ``finish()`` per view (cl.py:223-226); the
  result is bit-identical to the per-view loop.
* keyword-only ``device`` and ``i_range`` constructor arguments select the GPU and an axis-0
  slab of the grid (multi-GPU sharding, see ``parallel.py``).
* ``get_volume`` is an alias of ``get_values``; ``close()`` / context manager free the device
  memory deterministically.

There is no CPU fallback: without the built library or a CUDA device, construction raises.
"""
import ctypes
import logging
import os
import weakref
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import _cabi

logger = logging.getLogger(__name__)

EPS = 1e-10  # cl.py:35

_F = ctypes.POINTER(ctypes.c_float)
_I = ctypes.POINTER(ctypes.c_int32)

#: views handed to one bp_process_views call by process_label (host memory bound)
MAX_VIEWS_PER_CALL = 256
#: threads decoding mask images ahead of the device in process_label
DECODE_THREADS = min(16, os.cpu_count() or 1)


def img_as_float32(mask):
    """Host restatement of what ``skimage.util.img_as_float32`` does to the mask dtypes this
    path meets (used at cl.py:205-206): unsigned ints are scaled by ``1/max`` in float32,
    bool becomes 0/1, floats are cast."""
    mask = np.asarray(mask)
    if mask.dtype == np.float32:
        return mask
    if mask.dtype == np.bool_:
        return mask.astype(np.float32)
    if mask.dtype.kind == "u":
        return np.multiply(mask, 1.0 / np.iinfo(mask.dtype).max, dtype=np.float32)
    if mask.dtype.kind == "f":
        return mask.astype(np.float32)
    if mask.dtype.kind == "i":
        # skimage maps signed ints to [-1, 1]; masks never come this way in the pipeline
        info = np.iinfo(mask.dtype)
        out = np.multiply(mask, 1.0 / info.max, dtype=np.float32)
        return np.maximum(out, -1.0, out=out)
    raise TypeError(f"Unsupported mask dtype {mask.dtype}")


def _destroy_handle(h, keep):
    """Finaliser of a handle: the device work first, then the pinned buffers it may still be writing to."""
    _cabi.lib().bp_destroy(h)
    keep.clear()


class _PinnedArray:
    """numpy view over page-locked host memory from ``bp_host_alloc``."""

    def __init__(self, shape, dtype):
        n_bytes = int(np.prod(shape, dtype=np.int64)) * np.dtype(dtype).itemsize
        ptr = ctypes.c_void_p()
        _cabi.check(_cabi.lib().bp_host_alloc(ctypes.byref(ptr), n_bytes))
        self._ptr = ptr
        self._fin = weakref.finalize(self, _cabi.lib().bp_host_free, ptr)
        buf = (ctypes.c_char * max(n_bytes, 1)).from_address(ptr.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape, dtype=np.int64))).reshape(shape)
        # keep the allocation alive as long as any view of the array is
        self._keep = buf
        buf._owner = self


class Backprojection(object):
    """Back-projection onto a voxel volume (space carving or averaging).

    Same public surface as the reference class (cl.py:47-311).

    Parameters
    ----------
    shape : list
        Shape ``[nx, ny, nz]`` of the voxel volume.
    origin : list
        World position of voxel ``(0, 0, 0)``.
    voxel_size : float
        Size of a voxel.
    type : {'carving', 'averaging'}
        Kernel to use (cl.py:145-152); anything else raises ``ValueError``.
    default_value : float
        Seed of the grid (cl.py:173).
    labels : list, optional
        Labels of an ML pipeline mask set, or ``None``.
    log : bool
        For averaging, accumulate ``log(1e-10 + p)`` (cl.py:207-208).
    device : int, optional (keyword only)
        CUDA device ordinal; defaults to ``LOCAL_RANK`` (under torchrun) or 0.
    i_range : (int, int), optional (keyword only)
        Hold only planes ``[i0, i1)`` of axis 0 (slab sharding). ``get_values`` then returns
        ``[i1 - i0, ny, nz]``.
    speculative_readback : bool, optional (keyword only)
        Carving: let the library start copying the volume into ``values_h`` while host masks are still being uploaded
        (PCIe is full duplex; carved voxels are final), so that ``get_values`` only sends the rows that changed since.
        Same values as without it.  The one observable difference: ``values_h`` - the buffer a PREVIOUS ``get_values``
        returned a view of, as in the reference (cl.py:229-232) - may be rewritten during ``process_views`` /
        ``process_fileset`` instead of only inside the next ``get_values``; copy a result you want to keep across
        further views (the reference's own callers do, cl.py:249-254).  Default: on (``BP_SPEC_READBACK=0`` or
        ``False`` here switches it off).
    """

    def __init__(self, shape, origin, voxel_size, type="carving", default_value=0, labels=None, log=False,
                 *, device=None, i_range=None, stripes=None, speculative_readback=None):
        self.shape = shape
        self.origin = origin
        self.voxel_size = voxel_size
        self.default_value = default_value
        self.log = log
        self.labels = labels
        if type == "carving":
            self.dtype = np.int32
            self._mode = _cabi.MODE_CARVING
        elif type == "averaging":
            self.dtype = np.float32
            self._mode = _cabi.MODE_AVERAGING
        else:
            raise ValueError(f"Unknown kernel type {type}, valid values are 'averaging' or 'carving'!")
        self.kernel = type  # the reference stores the OpenCL kernel object here
        # Speculative read-back of the carve volume (see _register_readback_target): on unless switched off here or
        # with BP_SPEC_READBACK=0
        if speculative_readback is None:
            speculative_readback = os.environ.get("BP_SPEC_READBACK", "1") not in ("", "0")
        self.speculative_readback = bool(speculative_readback)
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        self.device = int(device)
        nx = int(self.shape[0])
        self.i_range = (0, nx) if i_range is None else (int(i_range[0]), int(i_range[1]))
        self.local_shape = [self.i_range[1] - self.i_range[0], int(self.shape[1]), int(self.shape[2])]
        # stripes = (lg_block, world, rank): hold the axis-0 planes whose block index (i >> lg_block) is
        # congruent to rank modulo world (load-balanced multi-GPU sharding, see parallel.py)
        self.stripes = None
        if stripes is not None:
            if i_range is not None:
                raise ValueError("give i_range or stripes, not both")
            lg, world, rank = (int(t) for t in stripes)
            self.stripes = (lg, world, rank)
            self.planes = np.array([i for i in range(nx) if (i >> lg) % world == rank], dtype=np.int64)
            self.local_shape[0] = len(self.planes)
        else:
            self.planes = np.arange(self.i_range[0], self.i_range[1], dtype=np.int64)

        buff_size = int(np.prod(self.local_shape, dtype=np.int64)) * 4
        logger.info(f"Buffer shape is {self.shape}")
        logger.info(f"Required memory for buffer is {buff_size / 2 ** 20:.1f} MiB!")

        self._h = None
        self._values = None
        self._lut_cache = {}
        self.init_buffers()

    # -- buffers ---------------------------------------------------------------------------
    def init_buffers(self):
        """Create the device grid seeded with ``default_value`` (cl.py:171-188)."""
        self.close()
        shape = (ctypes.c_int32 * 3)(*[int(s) for s in self.shape])
        origin = (ctypes.c_float * 3)(*[float(np.float32(o)) for o in self.origin])
        h = ctypes.c_void_p()
        if self.stripes is not None:
            _cabi.check(_cabi.lib().bp_create_striped(ctypes.byref(h), shape, *self.stripes, origin,
                                                      float(np.float32(self.voxel_size)), self._mode,
                                                      float(self.default_value), self.device))
        else:
            _cabi.check(_cabi.lib().bp_create(ctypes.byref(h), shape, self.i_range[0], self.i_range[1], origin,
                                              float(np.float32(self.voxel_size)), self._mode,
                                              float(self.default_value), self.device))
        self._h = h
        self._keep = []  # pinned buffers registered with the handle (speculative read-back) outlive it
        self._spec_registered = False
        self._fin = weakref.finalize(self, _destroy_handle, h, self._keep)
        return

    @property
    def values_h(self):
        """Host buffer ``get_values`` fills (pinned; allocated on first use)."""
        if self._values is None:
            self._values = _PinnedArray(tuple(self.local_shape), self.dtype)
            if self._values.array.size:
                self._values.array[...] = self.default_value
        return self._values.array

    def close(self):
        """Free the device memory now (also done by the garbage collector)."""
        if getattr(self, "_h", None) is not None:
            self._fin.detach()
            _cabi.lib().bp_destroy(self._h)
            self._h = None
            self._keep.clear()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _handle(self):
        if self._h is None:
            raise RuntimeError("Backprojection is closed")
        return self._h

    # -- mask conversion -------------------------------------------------------------------
    def _lut(self, kind):
        """256-entry table of what cl.py:205-208 makes of each byte value."""
        key = (kind, bool(self.log))
        if key not in self._lut_cache:
            if kind == "bool":
                lut = np.zeros(256, dtype=np.float32)
                lut[1:] = 1.0
            else:
                lut = img_as_float32(np.arange(256, dtype=np.uint8))
            if self.log:
                lut = np.log(EPS + lut)
            self._lut_cache[key] = np.ascontiguousarray(lut, dtype=np.float32)
        return self._lut_cache[key]

    def _prepare(self, mask, invert):
        """-> (contiguous array, mask_dtype code, flags, lut or None), applying on the host only
        what the device cannot do bit-identically."""
        mask = np.asarray(mask)
        if mask.ndim != 2:
            raise ValueError(f"mask must be a 2-D array, got shape {mask.shape}")
        flags = 0
        lut = None
        if self._mode == _cabi.MODE_CARVING:
            # cl.py:215 casts to int32 and the kernel tests `== 0`
            if mask.dtype == np.bool_:
                arr, code = mask.view(np.uint8), _cabi.MASK_U8
                flags |= _cabi.FLAG_BOOL_INVERT if invert else 0
            elif mask.dtype == np.uint8:
                arr, code = mask, _cabi.MASK_U8
                flags |= _cabi.FLAG_INVERT if invert else 0
            elif mask.dtype == np.uint16:
                arr, code = mask, _cabi.MASK_U16
                flags |= _cabi.FLAG_INVERT if invert else 0
            elif mask.dtype.kind in "iu":
                if invert:
                    mask = np.invert(mask)
                arr, code = mask.astype(np.int32), _cabi.MASK_I32
            elif mask.dtype.kind == "f":
                if invert:
                    np.invert(mask)  # raises TypeError exactly as the reference would
                arr, code = mask.astype(np.float32, copy=False), _cabi.MASK_F32
            else:
                raise TypeError(f"Unsupported mask dtype {mask.dtype}")
        else:
            if mask.dtype == np.bool_:
                arr, code, lut = mask.view(np.uint8), _cabi.MASK_U8, self._lut("bool")
                flags |= _cabi.FLAG_BOOL_INVERT if invert else 0
            elif mask.dtype == np.uint8:
                arr, code, lut = mask, _cabi.MASK_U8, self._lut("u8")
                flags |= _cabi.FLAG_INVERT if invert else 0
            else:
                if invert:
                    mask = np.invert(mask)
                m = img_as_float32(mask) if mask.dtype != np.float32 else mask  # cl.py:205-206
                if self.log:
                    m = np.log(EPS + m)  # cl.py:207-208
                arr, code = np.asarray(m, dtype=np.float32), _cabi.MASK_F32
        return np.ascontiguousarray(arr), code, flags, lut

    # -- processing ------------------------------------------------------------------------
    def process_view(self, intrinsics, rot, tvec, mask):
        """Process a new view (cl.py:190-227).

        Parameters
        ----------
        intrinsics: list
            [f_x, f_y, c_x, c_y]
        rot: list of list
            rotation matrix of the camera pose
        tvec: list
            translation vector of the camera pose
        mask: numpy.ndarray
            mask array (or float array if type is averaging)
        """
        self._process_batch([(intrinsics, rot, tvec, mask)], invert=False)
        return

    def process_views(self, views, invert=False):
        """Fused form: ``views`` is a sequence of ``(intrinsics, rot, tvec, mask)``; same result
        as calling ``process_view`` on each in order."""
        views = list(views)
        for s in range(0, len(views), MAX_VIEWS_PER_CALL):
            self._process_batch(views[s:s + MAX_VIEWS_PER_CALL], invert)

    def _register_readback_target(self):
        """Tell the library where ``get_values`` will want the carve volume, so that it can send an early copy while
        masks are still being uploaded (bp_set_readback_target; see ``speculative_readback``)."""
        if self._spec_registered or self._mode != _cabi.MODE_CARVING or not self.speculative_readback:
            return
        buf = self.values_h
        if buf.size:
            _cabi.check(_cabi.lib().bp_set_readback_target(self._handle(), ctypes.c_void_p(buf.ctypes.data)))
            self._keep.append(self._values)
        self._spec_registered = True

    def spec_snapshot(self):
        """Start the speculative read-back NOW: the carve volume as it stands begins its way to ``values_h`` while the
        caller goes on feeding views (bp_spec_snapshot; used by the multi-GPU sharded ingest, whose masks reach the
        library already packed, in chunks the library cannot foresee).  No-op unless ``speculative_readback``."""
        if self._mode != _cabi.MODE_CARVING or not self.speculative_readback:
            return
        self._register_readback_target()
        _cabi.check(_cabi.lib().bp_spec_snapshot(self._handle()))

    def spec_stats(self):
        """(rows written, rows compared) by the last speculative ``get_values`` (bp_spec_stats)."""
        a, b = ctypes.c_int64(), ctypes.c_int64()
        _cabi.check(_cabi.lib().bp_spec_stats(self._handle(), ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def _process_batch(self, views, invert):
        if not views:
            return
        self._register_readback_target()
        n = len(views)
        cams = np.empty((n, 16), dtype=np.float32)
        hw = np.empty((n, 2), dtype=np.int32)
        ptrs = (ctypes.c_void_p * n)()
        keep = []
        code0 = flags0 = lut0 = None
        groups = []  # consecutive views that share (dtype code, flags, lut) go in one call
        start = 0
        for q, (intr, rot, tvec, mask) in enumerate(views):
            cams[q, 0:4] = np.asarray(intr, dtype=np.float32).reshape(4)
            cams[q, 4:13] = np.asarray(rot, dtype=np.float32).reshape(9)
            cams[q, 13:16] = np.asarray(tvec, dtype=np.float32).reshape(3)
            arr, code, flags, lut = self._prepare(mask, invert)
            if logger.isEnabledFor(logging.DEBUG):
                logger.debug("mask max: %.2f" % (arr.max()))  # cl.py:214
            keep.append(arr)
            hw[q] = arr.shape
            ptrs[q] = arr.ctypes.data
            sig = (code, flags, id(lut))
            if q == 0:
                code0, flags0, lut0 = code, flags, lut
                sig0 = sig
            elif sig != sig0:
                groups.append((start, q, code0, flags0, lut0))
                start, code0, flags0, lut0, sig0 = q, code, flags, lut, sig
        groups.append((start, n, code0, flags0, lut0))
        L = _cabi.lib()
        for (a, b, code, flags, lut) in groups:
            lut_p = lut.ctypes.data_as(_F) if lut is not None else None
            sub_ptrs = (ctypes.c_void_p * (b - a))(*ptrs[a:b])
            _cabi.check(L.bp_process_views(self._handle(), b - a, cams[a:b].ctypes.data_as(_F), sub_ptrs,
                                           hw[a:b].ctypes.data_as(_I), code, flags, lut_p))

    def process_device_views(self, cams, mask_ptrs, hw, invert=False, dtype="uint8"):
        """Fused pass over views whose masks already sit in device memory (inputs resident in HBM):
        ``cams`` float32 ``[n, 16]`` (fx fy cx cy | rot | tvec), ``mask_ptrs`` n device addresses of
        row-major ``H x W`` arrays of ``dtype`` ('uint8' or 'float32'), ``hw`` int ``[n, 2]``."""
        n = len(mask_ptrs)
        cams = np.ascontiguousarray(cams, dtype=np.float32).reshape(n, 16)
        hw = np.ascontiguousarray(hw, dtype=np.int32).reshape(n, 2)
        if dtype == "uint8":
            code, flags = _cabi.MASK_U8, (_cabi.FLAG_INVERT if invert else 0)
            lut = self._lut("u8") if self._mode == _cabi.MODE_AVERAGING else None
        elif dtype == "float32":
            if invert:
                raise TypeError("invert is not defined for float masks")
            code, flags, lut = _cabi.MASK_F32, 0, None
        else:
            raise TypeError(f"Unsupported device mask dtype {dtype}")
        flags |= _cabi.FLAG_DEVICE_MASKS
        L = _cabi.lib()
        for a in range(0, n, MAX_VIEWS_PER_CALL):
            b = min(n, a + MAX_VIEWS_PER_CALL)
            ptrs = (ctypes.c_void_p * (b - a))(*[int(p) for p in mask_ptrs[a:b]])
            _cabi.check(L.bp_process_views(self._handle(), b - a, cams[a:b].ctypes.data_as(_F), ptrs,
                                           hw[a:b].ctypes.data_as(_I), code, flags,
                                           lut.ctypes.data_as(_F) if lut is not None else None))

    # -- sharded ingest (multi-GPU slab mode, see parallel.py) ------------------------------------
    @staticmethod
    def packed_size(H, W):
        """Bytes one ``H x W`` mask packs to (32x32-pixel bit tiles)."""
        n = ctypes.c_int64()
        _cabi.check(_cabi.lib().bp_packed_size(int(H), int(W), ctypes.byref(n)))
        return n.value

    def pack_views(self, masks, dst_ptr, view_stride, invert=False, device_masks=False, hw=None, cams=None):
        """Upload + bit-pack ``masks`` (2-D uint8 / bool arrays, or with ``device_masks`` device addresses of
        uint8 arrays whose sizes are ``hw``) to ``dst_ptr + q * view_stride`` in device memory.  Carving only.
        ``cams`` (float32 ``[n, 16]``, optional): lets the library trim each view to the rectangle the whole grid
        projects into (bp_pack_views_cams; ``BP_VIEW_RECT=0`` switches the trimming off)."""
        n = len(masks)
        if n == 0:
            return
        ptrs = (ctypes.c_void_p * n)()
        keep = []
        if device_masks:
            hw = np.ascontiguousarray(hw, dtype=np.int32).reshape(n, 2)
            for q in range(n):
                ptrs[q] = int(masks[q])
            code, flags = _cabi.MASK_U8, (_cabi.FLAG_INVERT if invert else 0) | _cabi.FLAG_DEVICE_MASKS
        else:
            hw = np.empty((n, 2), dtype=np.int32)
            code = flags = None
            for q, m in enumerate(masks):
                arr, c, f, _ = self._prepare(m, invert)
                if code is None:
                    code, flags = c, f
                elif (c, f) != (code, flags):
                    raise TypeError("pack_views needs masks of one dtype")
                keep.append(arr)
                hw[q] = arr.shape
                ptrs[q] = arr.ctypes.data
        if cams is not None:
            cams = np.ascontiguousarray(cams, dtype=np.float32).reshape(n, 16)
            _cabi.check(_cabi.lib().bp_pack_views_cams(self._handle(), n, cams.ctypes.data_as(_F), ptrs,
                                                       hw.ctypes.data_as(_I), code, flags,
                                                       ctypes.c_void_p(int(dst_ptr)), int(view_stride)))
            return
        _cabi.check(_cabi.lib().bp_pack_views(self._handle(), n, ptrs, hw.ctypes.data_as(_I), code, flags,
                                              ctypes.c_void_p(int(dst_ptr)), int(view_stride)))

    def process_packed_views(self, cams, packed_ptrs, hw):
        """Fused carve over views whose bit tiles (``pack_views`` layout) already sit in device memory."""
        n = len(packed_ptrs)
        cams = np.ascontiguousarray(cams, dtype=np.float32).reshape(n, 16)
        hw = np.ascontiguousarray(hw, dtype=np.int32).reshape(n, 2)
        L = _cabi.lib()
        for a in range(0, n, MAX_VIEWS_PER_CALL):
            b = min(n, a + MAX_VIEWS_PER_CALL)
            ptrs = (ctypes.c_void_p * (b - a))(*[int(p) for p in packed_ptrs[a:b]])
            _cabi.check(L.bp_process_views(self._handle(), b - a, cams[a:b].ctypes.data_as(_F), ptrs,
                                           hw[a:b].ctypes.data_as(_I), _cabi.MASK_U8, _cabi.FLAG_PACKED_MASKS, None))

    def get_values(self):
        """Gets computed values from the device (cl.py:229-232)."""
        out = self.values_h
        _cabi.check(_cabi.lib().bp_get_values(self._handle(), out.ctypes.data_as(ctypes.c_void_p)))
        return out.reshape(self.local_shape)

    get_volume = get_values  # the name BASELINE.json's north_star uses

    def get_values_tail(self, out=None, float64=False, exp_clip=False):
        """``get_values`` with the tail of ``Voxels.run`` fused into the read-back (tasks/cl.py:166-174):
        returns ``(values, vmin, vmax)`` - ``vmin`` / ``vmax`` of the raw values replace the ``np.unique`` sanity
        check (tasks/cl.py:168); ``float64`` writes the dtype ``process_fileset`` stacks labels into (cl.py:249);
        ``exp_clip`` applies ``np.exp`` and the clip to 1 of tasks/cl.py:172-174 (averaging only).  ``out`` may be
        any C-contiguous array of the right size and dtype (e.g. a row of the stacked result); default: a new array
        (float64) or the pinned ``values_h``."""
        dt = np.dtype(np.float64) if float64 else self.dtype
        if out is None:
            out = np.empty(tuple(self.local_shape), dtype=dt) if float64 else self.values_h
        if out.dtype != dt or not out.flags["C_CONTIGUOUS"] or out.size != int(np.prod(self.local_shape)):
            raise ValueError(f"out must be a C-contiguous {dt} array of {int(np.prod(self.local_shape))} elements")
        lo, hi = ctypes.c_double(), ctypes.c_double()
        flags = (_cabi.TAIL_F64 if float64 else 0) | (_cabi.TAIL_EXP_CLIP if exp_clip else 0)
        _cabi.check(_cabi.lib().bp_get_values_tail(self._handle(), out.ctypes.data_as(ctypes.c_void_p), flags,
                                                   ctypes.byref(lo), ctypes.byref(hi)))
        return out.reshape(self.local_shape), lo.value, hi.value

    def process_fileset(self, fs, camera_metadata, invert=False, exp_clip=False, stats=None):
        """Processes a whole fileset (cl.py:234-257).

        Parameters
        ----------
        fs : plantdb.db.Fileset or list of plantdb.db.File
            The images `Fileset` or list of images `File` to process.
        camera_metadata : str
            Name of the metadata to use to get the camera intrinsics (fx, fy, cx, cy) & poses.
        invert : bool, optional
            If ``True``, invert the values of the mask file to process.
        exp_clip : bool, optional
            Not in the reference: fuse the ``np.exp`` + clip of ``Voxels.run`` (tasks/cl.py:172-174) into the
            read-back.  ``stats``, a dict, receives ``min`` / ``max`` of the raw volume (tasks/cl.py:168).
        """
        lo_all, hi_all = np.inf, -np.inf
        if self.labels is not None:
            result = np.zeros((len(self.labels), *self.local_shape))
            for i, label in enumerate(self.labels):
                logger.info(f"Processing label '{label}'...")
                if i != 0:
                    self.clear()
                self.process_label(fs, camera_metadata, label, invert, fetch=False)
                # float64 conversion (exact), min / max and the optional exp + clip happen on the device
                _, lo, hi = self.get_values_tail(out=result[i], float64=True, exp_clip=exp_clip)
                lo_all, hi_all = min(lo_all, lo), max(hi_all, hi)  # NaN propagates like numpy's min / max
                if lo != lo:
                    lo_all = hi_all = np.nan
        else:
            self.process_label(fs, camera_metadata, None, invert=invert, fetch=False)
            result, lo_all, hi_all = self.get_values_tail(exp_clip=exp_clip)
        if stats is not None:
            stats["min"], stats["max"] = lo_all, hi_all
        return result

    def process_label(self, fs, camera_metadata, label=None, invert=False, fetch=True):
        """Processes a whole fileset for given label (cl.py:259-305).

        Returns
        -------
        numpy.ndarray
            The processed volume, for given label, if any.
        """
        if hasattr(fs, "get_files"):
            fs = fs.get_files()

        todo = []
        for fi in fs:
            # Skip file if not of the right label (when defined)
            if label is not None and fi.get_metadata("channel") != label:
                continue
            logger.debug("processing file %s" % fi.id)
            cam = fi.get_metadata(camera_metadata, default=None)
            if cam is None:
                logger.warning(f"Could not get camera params from '{camera_metadata}' for {fi.id}, skipping...")
                continue
            intrinsics = np.array(cam["camera_model"]['params'][0:4], dtype=np.float32)
            rot = np.array(sum(cam['rotmat'], []), dtype=np.float32)
            tvec = np.array(cam['tvec'], dtype=np.float32)
            todo.append((intrinsics, rot, tvec, fi))
        # Mask ingest: the image decode of cl.py:297 runs on a thread pool (PIL / imageio release the
        # GIL while decoding), one launch-sized batch ahead of the device, in file order.
        with ThreadPoolExecutor(max_workers=min(DECODE_THREADS, max(len(todo), 1))) as pool:
            chunks = [todo[o:o + MAX_VIEWS_PER_CALL] for o in range(0, len(todo), MAX_VIEWS_PER_CALL)]
            pending = [pool.submit(_read_image, t[3]) for t in chunks[0]] if chunks else []
            for c, chunk in enumerate(chunks):
                masks = [p.result() for p in pending]
                pending = [pool.submit(_read_image, t[3]) for t in chunks[c + 1]] if c + 1 < len(chunks) else []
                self._process_batch([(t[0], t[1], t[2], m) for t, m in zip(chunk, masks)], invert)
        if not chunks:
            self._process_batch([], invert)
        return self.get_values() if fetch else None

    def clear(self):
        """Re-seed the device grid with ``default_value`` (cl.py:307-311)."""
        _cabi.check(_cabi.lib().bp_clear(self._handle()))
        return

    # -- extras used by parallel.py / bench.py ------------------------------------------------
    def sync(self):
        _cabi.check(_cabi.lib().bp_sync(self._handle()))

    def stream(self):
        """The handle's cudaStream_t as an int (for torch.cuda.ExternalStream)."""
        p = ctypes.c_void_p()
        _cabi.check(_cabi.lib().bp_stream(self._handle(), ctypes.byref(p)))
        return p.value or 0

    def timing(self):
        """(kernel_ms, ingest_ms, launches): device time since the last call, launches since creation."""
        k, g, n = ctypes.c_float(), ctypes.c_float(), ctypes.c_int64()
        _cabi.check(_cabi.lib().bp_timing(self._handle(), ctypes.byref(k), ctypes.byref(g), ctypes.byref(n)))
        return k.value, g.value, n.value

    def work_stats(self):
        """(projections evaluated, cull blocks, cull blocks written off as carved): diagnostics of the carve path
        (bp_work_stats; the first is counted only under ``BP_COUNT_WORK=1``)."""
        a, b, c = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64()
        _cabi.check(_cabi.lib().bp_work_stats(self._handle(), ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return a.value, b.value, c.value

    def ingest_bytes(self):
        """Mask bytes copied host -> device on this handle since creation (see bp_ingest_bytes)."""
        n = ctypes.c_int64()
        _cabi.check(_cabi.lib().bp_ingest_bytes(self._handle(), ctypes.byref(n)))
        return n.value

    @staticmethod
    def view_rect(shape, origin, voxel_size, cam, H, W, planes=None):
        """``(x0, y0, x1, y1)``: the pixels of a view that the voxels of the grid (or of its axis-0 planes
        ``planes = (begin, end)``) can project to (see bp_view_rect; host only, needs no GPU).  ``cam`` is the 16-float
        row ``fx fy cx cy | rot | tvec``."""
        shp = np.asarray(shape, dtype=np.int32)
        b, e = (0, int(shp[0])) if planes is None else (int(planes[0]), int(planes[1]))
        org = np.ascontiguousarray(origin, dtype=np.float32)
        c = np.ascontiguousarray(cam, dtype=np.float32).reshape(16)
        rect = np.zeros(4, dtype=np.int32)
        _cabi.check(_cabi.lib().bp_view_rect(shp.ctypes.data_as(_I), b, e, org.ctypes.data_as(_F), float(voxel_size),
                                             c.ctypes.data_as(_F), int(H), int(W), rect.ctypes.data_as(_I)))
        return tuple(int(v) for v in rect)

    def device_state(self):
        """(device pointer, bytes, nz pitch) of the raw state (see bp_device_state)."""
        p, nb, pitch = ctypes.c_void_p(), ctypes.c_int64(), ctypes.c_int32()
        _cabi.check(_cabi.lib().bp_device_state(self._handle(), ctypes.byref(p), ctypes.byref(nb),
                                                ctypes.byref(pitch)))
        return p.value, nb.value, pitch.value


    def state_tensor(self):
        """torch tensor ALIASING the raw device state ``[planes, ny, nz_pitch]`` (int8 carving coding
        0 carved < 1 seen < 2 unseen, or float32 partial sums): what ``parallel.py`` hands to NCCL."""
        import torch
        ptr, n_bytes, pitch = self.device_state()
        carving = self._mode == _cabi.MODE_CARVING
        shape = (self.local_shape[0], self.local_shape[1], pitch)

        class _Span:  # minimal __cuda_array_interface__ exporter over memory owned by the handle
            __cuda_array_interface__ = {"shape": shape, "typestr": "|i1" if carving else "<f4",
                                        "data": (ptr or 0, False), "version": 2}

        if ptr is None or n_bytes == 0:
            return torch.empty(shape, dtype=torch.int8 if carving else torch.float32, device=f"cuda:{self.device}")
        t = torch.as_tensor(_Span(), device=f"cuda:{self.device}")
        t._bp_owner = self  # keep the handle alive as long as the alias is
        return t

    def adopt_state(self, state):
        """Called after an in-place collective on ``state_tensor()`` ran on torch's current stream:
        make the result visible to the handle's own stream."""
        import torch
        torch.cuda.current_stream(self.device).synchronize()


def _read_image(fi):
    """``plantdb.io.read_image`` (io.py:343-378) for a plantdb ``File`` when plantdb is
    installed, else for the duck-typed files of ``fsdb_lite`` (``read_image()`` / ``path()``)."""
    if hasattr(fi, "read_image"):
        return fi.read_image()
    try:
        from plantdb import io
        return io.read_image(fi)
    except ImportError:
        from PIL import Image
        path = fi.path() if callable(getattr(fi, "path", None)) else fi
        return np.asarray(Image.open(path))

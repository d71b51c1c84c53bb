#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.proc3d
_________________________

Volume -> point-cloud handoff on the GPU: drop-in for ``proc3d.vol2pcd`` of the
reference (plant3dvision/proc3d.py:490-570) and for the multiclass gates of its ``PointCloud`` task
(plant3dvision/tasks/proc3d.py:66-136).  Everything numeric runs in ``libbp_b200.so``
(``csrc/vol2pcd_b200.cu`` behind ``include/vol2pcd_b200.h``); there is no CPU fallback.

``vol2pcd(volume, origin, voxel_size, level_set_value)`` has the reference's signature.  The
reference returns an ``open3d.geometry.PointCloud``; Open3D is not a dependency here, so the result is
a small :class:`PointCloud` holding ``points`` / ``normals`` as float64 ``[N, 3]`` arrays (what
``np.asarray(pcd.points)`` / ``np.asarray(pcd.normals)`` give) with ``to_open3d()`` for callers that
have Open3D.  ``volume`` may also be a live :class:`~plant3dvision_b200.cl.Backprojection`: the state
is then read where it lies in HBM (no D2H / TIFF / H2D round trip, tasks/cl.py:176-184 ->
tasks/proc3d.py:69-78).
"""
import ctypes
import logging

import numpy as np

from . import _cabi

logger = logging.getLogger(__name__)

_SQRT3 = np.sqrt(3)

_DTYPES = {np.dtype(np.uint8): _cabi.V2P_U8, np.dtype(np.bool_): _cabi.V2P_U8, np.dtype(np.int8): _cabi.V2P_I8,
           np.dtype(np.int32): _cabi.V2P_I32, np.dtype(np.float32): _cabi.V2P_F32,
           np.dtype(np.float64): _cabi.V2P_F64}


def gaussian_weights(sigma=1.0, truncate=4.0):
    """The weights ``scipy.ndimage.gaussian_filter(., sigma)`` correlates with (scipy
    ``_gaussian_kernel1d``, order 0): radius ``int(truncate * sigma + 0.5)``, ``exp(-x^2 / (2 sigma^2))``
    normalised by its sum.  Computed with numpy so that they are bit-identical to scipy's."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi_x = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    phi_x = phi_x / phi_x.sum()
    return np.ascontiguousarray(phi_x, dtype=np.float64), radius


class PointCloud:
    """Points and unit normals, float64 ``[N, 3]`` (the two attributes the reference fills at
    proc3d.py:565-568)."""

    def __init__(self, points, normals, timing=None):
        self.points = points
        self.normals = normals
        self.timing = timing or {}

    def __len__(self):
        return len(self.points)

    def to_open3d(self):
        import open3d as o3d
        pcd = o3d.geometry.PointCloud()
        pcd.points = o3d.utility.Vector3dVector(self.points)
        pcd.normals = o3d.utility.Vector3dVector(self.normals)
        return pcd


def index2point(indexes, origin, voxel_size):
    """proc3d.py:28-45."""
    return voxel_size * indexes + np.asarray(origin)[np.newaxis, :]


def point2index(points, origin, voxel_size):
    """proc3d.py:48-66."""
    return np.array(np.round((points - np.asarray(origin)[np.newaxis, :]) / voxel_size), dtype=int)


def _log_cutoff(dtype):
    """Smallest value v of ``dtype`` with ``np.exp(v) > 0.5`` (np.exp is monotone): thresholding a
    log-domain sum at ``>= cutoff`` selects exactly the voxels the reference keeps after the
    ``np.exp`` of tasks/cl.py:172-174 and the ``> 0.5`` of proc3d.py:515."""
    dt = np.dtype(dtype)
    v = dt.type(np.log(0.5))
    while np.exp(v) > 0.5:
        v = np.nextafter(v, dt.type(-np.inf))
    while not (np.exp(v) > 0.5):
        v = np.nextafter(v, dt.type(np.inf))
    return float(v)


def _run(vol_ptr, code, flags, shape, pitch, bin_op, bin_value, origin, voxel_size, level_set_value, device,
         return_distance=False):
    L = _cabi.lib()
    w, radius = gaussian_weights(1.0)
    res = ctypes.c_void_p()
    shp = (ctypes.c_int32 * 3)(*[int(s) for s in shape])
    org = np.ascontiguousarray(np.asarray(origin, dtype=np.float64).reshape(3))
    lsv = float(level_set_value)
    band_hi = float(-level_set_value + _SQRT3)   # proc3d.py:536
    half_diag = float(_SQRT3 / 2)                 # proc3d.py:547
    D = ctypes.POINTER(ctypes.c_double)
    _cabi.check(L.bp_vol2pcd(ctypes.byref(res), int(device), ctypes.c_void_p(vol_ptr), code, flags, shp, int(pitch),
                             bin_op, float(bin_value), org.ctypes.data_as(D), float(voxel_size), lsv, band_hi,
                             half_diag, w.ctypes.data_as(D), radius))
    try:
        n, tot, edt, band, launches = (ctypes.c_int64(), ctypes.c_float(), ctypes.c_float(), ctypes.c_int64(),
                                       ctypes.c_int64())
        _cabi.check(L.bp_vol2pcd_info(res, ctypes.byref(n), ctypes.byref(tot), ctypes.byref(edt), ctypes.byref(band),
                                      ctypes.byref(launches)))
        pts = np.empty((n.value, 3), dtype=np.float64)
        nrm = np.empty((n.value, 3), dtype=np.float64)
        _cabi.check(L.bp_vol2pcd_fetch(res, pts.ctypes.data_as(ctypes.c_void_p), nrm.ctypes.data_as(ctypes.c_void_p)))
        timing = {"total_ms": tot.value, "edt_ms": edt.value, "band_voxels": band.value, "launches": launches.value}
        out = PointCloud(pts, nrm, timing)
        if return_distance:
            dist = np.empty(tuple(int(s) for s in shape), dtype=np.float64)
            cap = ctypes.c_double()
            _cabi.check(L.bp_vol2pcd_signed_distance(res, dist.ctypes.data_as(ctypes.c_void_p), ctypes.byref(cap)))
            out.signed_distance, out.distance_cap = dist, cap.value
        return out
    finally:
        L.bp_vol2pcd_free(res)


def vol2pcd(volume, origin, voxel_size, level_set_value=0, *, device=None, return_distance=False):
    """Converts a volume into a point-cloud with normals (proc3d.py:490-570).

    ``volume``: ``NxMxP`` numpy array (bool / uint8 / int8 / int32 / float32 / float64; other dtypes are
    cast to float64 like the reference's ``volume > 0.5`` would promote them), or a live
    ``Backprojection`` whose device state is used in place."""
    from .cl import Backprojection
    if isinstance(volume, Backprojection):
        return _vol2pcd_from_backprojection(volume, origin, voxel_size, level_set_value, return_distance)
    if device is None:
        import os
        device = int(os.environ.get("LOCAL_RANK", "0"))
    volume = np.asarray(volume)
    if volume.ndim != 3:
        raise ValueError(f"vol2pcd needs a 3-D volume, got shape {volume.shape}")
    if volume.dtype not in _DTYPES:
        volume = volume.astype(np.float64)
    volume = np.ascontiguousarray(volume)
    logger.info("Volume binarization, distance transform, gradient, Gaussian filtering, point detection on the GPU...")
    return _run(volume.ctypes.data, _DTYPES[volume.dtype], 0, volume.shape, 0, _cabi.V2P_BIN_GT, 0.5, origin,
                voxel_size, level_set_value, device, return_distance)


def _vol2pcd_from_backprojection(bp, origin, voxel_size, level_set_value, return_distance=False):
    """Fused handoff: binarise the handle's raw device state exactly as the reference binarises
    what ``Voxels`` would have written (tasks/cl.py:172-184 then proc3d.py:515)."""
    if getattr(bp, "stripes", None) is not None:
        raise ValueError("vol2pcd needs contiguous planes: gather a striped volume first")
    ptr, _, pitch = bp.device_state()
    bp.sync()
    shape = bp.local_shape
    if origin is None:
        origin = [float(bp.origin[0]) + bp.i_range[0] * float(bp.voxel_size), float(bp.origin[1]), float(bp.origin[2])]
    if bp._mode == _cabi.MODE_CARVING:
        # int8 state 0 carved / 1 seen / 2 unseen <-> values -1 / (1 if default == 0 else default) / default
        d = int(bp.default_value)
        if d == -1 or not (d > 0.5 or d == 0):
            op, val = _cabi.V2P_BIN_GT, 3.0     # nothing exceeds 0.5: empty foreground
        elif d == 0:
            op, val = _cabi.V2P_BIN_EQ, 1.0     # only 'seen' voxels hold 1
        else:
            op, val = _cabi.V2P_BIN_GE, 1.0     # seen and unseen both hold default > 0.5
        code = _cabi.V2P_I8
    else:
        code = _cabi.V2P_F32
        if bp.log:
            op, val = _cabi.V2P_BIN_GE, _log_cutoff(np.float32)  # exp(v) > 0.5 (clip at 1 is irrelevant)
        else:
            op, val = _cabi.V2P_BIN_GT, 0.5
    return _run(ptr, code, _cabi.V2P_FLAG_DEVICE_VOLUME, shape, pitch, op, val, origin, voxel_size, level_set_value,
                bp.device, return_distance)


def multiclass_select(voxels, background_prior=1.0, min_contrast=10.0, min_score=0.2, *, device=None):
    """tasks/proc3d.py:80-115 on the GPU.  ``voxels``: mapping label -> volume (the Voxels NPZ).
    Yields ``(label, device_pointer, shape)`` for every non-background label; the pointer holds the uint8
    0/1 volume ``pred_c`` and is freed when the generator advances."""
    L = _cabi.lib()
    if device is None:
        import os
        device = int(os.environ.get("LOCAL_RANK", "0"))
    labels = list(voxels.keys())
    vols = [np.asarray(voxels[l]) for l in labels]
    dt = np.float32 if all(v.dtype == np.float32 for v in vols) else np.float64
    vols = [np.ascontiguousarray(v, dtype=dt) for v in vols]
    shape = vols[0].shape
    if any(v.shape != shape for v in vols):
        raise ValueError("label volumes differ in shape")
    bg = labels.index("background") if "background" in labels else -1
    ptrs = (ctypes.c_void_p * len(vols))(*[v.ctypes.data for v in vols])
    shp = (ctypes.c_int32 * 3)(*[int(s) for s in shape])
    code = _cabi.V2P_F32 if dt == np.float32 else _cabi.V2P_F64
    for i, l in enumerate(labels):
        if l == "background":
            continue
        out = ctypes.c_void_p()
        _cabi.check(L.bp_multiclass_select(int(device), len(vols), ptrs, code, 0, shp, bg, float(background_prior),
                                           float(min_contrast), float(min_score), i, ctypes.byref(out)))
        try:
            yield l, out.value, shape
        finally:
            L.bp_device_free(int(device), out)


def pointcloud_from_voxels(voxels, origin, voxel_size, level_set_value=1.0, background_prior=1.0, min_contrast=10.0,
                           min_score=0.2, *, device=None):
    """``PointCloud.run`` (tasks/proc3d.py:66-136) without the file I/O: ``voxels`` is a single volume
    (TIFF / single-key NPZ case) or a mapping label -> volume (multiclass).  Returns
    ``(PointCloud, point_labels or None)``; defaults are the task's (tasks/proc3d.py:60-64)."""
    if device is None:
        import os
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if isinstance(voxels, dict) and len(voxels) == 1:
        voxels = voxels[list(voxels.keys())[0]]
    if not isinstance(voxels, dict):
        return vol2pcd(voxels, np.asarray(origin), float(voxel_size), level_set_value, device=device), None
    pts, nrm, point_labels = [np.zeros((0, 3))], [np.zeros((0, 3))], []
    for label, ptr, shape in multiclass_select(voxels, background_prior, min_contrast, min_score, device=device):
        out = _run(ptr, _cabi.V2P_U8, _cabi.V2P_FLAG_DEVICE_VOLUME, shape, 0, _cabi.V2P_BIN_GT, 0.5, origin,
                   float(voxel_size), level_set_value, device)
        pts.append(out.points)
        nrm.append(out.normals)
        point_labels += [label] * len(out.points)
    return PointCloud(np.concatenate(pts), np.concatenate(nrm)), point_labels


# ---------------------------------------------------------------------------------------------------
# SegmentedPointCloud scoring (plant3dvision/tasks/proc3d.py:185-253)
# ---------------------------------------------------------------------------------------------------
def backproject_points(points, K, rot, tvec):
    """proc3d.py:655-659, verbatim (host numpy; the GPU path below does the same arithmetic per point)."""
    x = rot @ points.transpose() + tvec[:, np.newaxis]
    x = K @ x
    x = x / x[2, :][np.newaxis, :]
    return x[:2, :].transpose()


def score_points(points, views, labels, *, device=None, return_scores=True):
    """The scoring loops of ``SegmentedPointCloud.run`` (tasks/proc3d.py:203-230) on the GPU.

    ``points``: ``[N, 3]`` float64.  ``views``: sequence of ``(label, camera, mask)`` in file order, ``camera`` the
    metadata dict the task reads (``rotmat``, ``tvec``, ``camera_model.params``), ``mask`` a 2-D uint8 label map;
    views whose label is not in ``labels`` or whose camera is ``None`` are skipped like the reference does
    (tasks/proc3d.py:205-217).  ``labels``: the label list IN THE ORDER the caller wants ties broken (the reference
    builds it from a Python ``set``, so its order is arbitrary; tasks/proc3d.py:193-201).
    Returns ``(pts_labels [N] int, scores [L, N] float64 or None, device_ms)``."""
    L = _cabi.lib()
    if device is None:
        import os
        device = int(os.environ.get("LOCAL_RANK", "0"))
    pts = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
    labels = list(labels)
    cams, maps, hw, vlab = [], [], [], []
    for label, camera, mask in views:
        if label not in labels:
            continue
        if camera is None:
            logger.warning("Could not get camera pose for view, skipping...")
            continue
        m = np.asarray(mask)
        if m.ndim != 2 or m.dtype not in (np.uint8, np.bool_):
            raise TypeError(f"label maps must be 2-D uint8 arrays, got {m.dtype} {m.shape}")
        m = np.ascontiguousarray(m.view(np.uint8))
        intr = camera["camera_model"]["params"]
        cams.append(np.concatenate([np.asarray(intr[0:4], dtype=np.float64),
                                    np.asarray(camera["rotmat"], dtype=np.float64).reshape(9),
                                    np.asarray(camera["tvec"], dtype=np.float64).reshape(3)]))
        maps.append(m)
        hw.append(m.shape)
        vlab.append(labels.index(label))
    n, nv, nl = len(pts), len(maps), max(len(labels), 1)
    cams_a = np.ascontiguousarray(np.array(cams, dtype=np.float64).reshape(nv, 16))
    hw_a = np.ascontiguousarray(np.array(hw, dtype=np.int32).reshape(nv, 2))
    vlab_a = np.ascontiguousarray(np.array(vlab, dtype=np.int32))
    ptrs = (ctypes.c_void_p * max(nv, 1))(*[m.ctypes.data for m in maps])
    scores = np.zeros((nl, n), dtype=np.float64) if return_scores else None
    out = np.zeros(n, dtype=np.int32)
    ms = ctypes.c_float()
    D, I = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int32)
    _cabi.check(L.bp_score_points(int(device), n, pts.ctypes.data_as(D), nv, cams_a.ctypes.data_as(D), ptrs,
                                  hw_a.ctypes.data_as(I), vlab_a.ctypes.data_as(I), nl,
                                  scores.ctypes.data_as(D) if scores is not None else None, out.ctypes.data_as(I),
                                  ctypes.byref(ms)))
    return out.astype(np.int64), scores, ms.value

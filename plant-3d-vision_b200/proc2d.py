#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.proc2d
_________________________

The mask producer on the GPU: what ``plant3dvision.tasks.proc2d.Masks.f`` does per image
(tasks/proc2d.py:224-237) with ``plant3dvision.proc2d.linear`` / ``excess_green`` / ``dilation``
(proc2d.py:69-220), for a whole batch of images in one call of ``bp_masks_filter`` (``include/masks_b200.h``).
There is no CPU fallback: without the CUDA library these functions raise.
"""
import ctypes
import logging
import os

import numpy as np

from . import _cabi

logger = logging.getLogger(__name__)

CROSS = np.array([[0, 1, 0], [1, 1, 1], [0, 1, 0]], dtype=np.uint8)


def dilation_sequence(n):
    """The footprint sequence ``proc2d.dilation(img, n)`` uses (proc2d.py:219:
    ``disk(n, decomposition='sequence')``) as ``((3x3 footprint, repeats), ...)``.  The decomposition table lives
    inside scikit-image, so radii > 1 need scikit-image importable (it is wherever the reference runs); radius 1
    is the exact cross; 0 is no dilation."""
    n = int(n)
    if n <= 0:
        return ()
    try:
        from skimage.morphology import disk
    except ImportError:
        if n == 1:
            return ((CROSS, 1),)
        raise ImportError("dilation > 1 needs skimage.morphology.disk(n, decomposition='sequence'); "
                          "pass the footprint sequence explicitly instead") from None
    return tuple((np.asarray(fp, dtype=np.uint8), int(k)) for fp, k in disk(n, decomposition="sequence"))


def masks(images, type="linear", parameters=(0, 1, 0), threshold=0.3, dilation=0, *, device=None,
          keep_on_device=False):
    """Binary masks of a batch of RGB(A) uint8 images of one size -> list of uint8 ``[H, W]`` arrays (0 / 255).

    ``type``, ``parameters``, ``threshold``, ``dilation``: the parameters of the reference task (tasks/proc2d.py:207-211);
    ``dilation`` may also be an explicit footprint sequence ``((3x3 array, repeats), ...)``.
    ``keep_on_device=True`` returns ``(device_pointer, n, H, W, device_ms)`` instead: the n masks back to back in
    one device buffer (free it with ``device_free``), ready for ``Backprojection.process_device_views``."""
    L = _cabi.lib()
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if type == "linear":
        t = _cabi.MASKS_LINEAR
        coefs = (ctypes.c_double * 3)(*[float(c) for c in list(parameters)])
    elif type == "excess_green":
        t, coefs = _cabi.MASKS_EXCESS_GREEN, None
    else:
        raise Exception(f"Unknown masking type '{type}'!")  # tasks/proc2d.py:221-222
    imgs = [np.ascontiguousarray(im) for im in images]
    n = len(imgs)
    if n == 0:
        return (None, 0, 0, 0, 0.0) if keep_on_device else []
    H, W = imgs[0].shape[:2]
    for im in imgs:
        if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] not in (3, 4):
            raise TypeError(f"images must be uint8 [H, W, 3|4] arrays, got {im.dtype} {im.shape}")
        if im.shape != imgs[0].shape:
            raise ValueError("all images of one call must have the same shape")
    C = imgs[0].shape[2]
    seq = dilation_sequence(dilation) if np.isscalar(dilation) else tuple(dilation)
    fps = np.zeros((max(len(seq), 1), 9), dtype=np.uint8)
    reps = np.zeros(max(len(seq), 1), dtype=np.int32)
    for f, (fp, k) in enumerate(seq):
        fp = np.asarray(fp)
        if fp.shape != (3, 3):
            raise ValueError(f"footprints must be 3x3, got {fp.shape}")
        fps[f] = (fp != 0).reshape(9)
        reps[f] = int(k)
    vp = ctypes.c_void_p
    in_ptrs = (vp * n)(*[im.ctypes.data for im in imgs])
    outs = None if keep_on_device else [np.empty((H, W), dtype=np.uint8) for _ in range(n)]
    out_ptrs = None if keep_on_device else (vp * n)(*[o.ctypes.data for o in outs])
    dev = vp()
    ms = ctypes.c_float()
    _cabi.check(L.bp_masks_filter(int(device), n, in_ptrs, H, W, C, t, coefs, float(threshold), len(seq),
                                  fps.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)),
                                  reps.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), out_ptrs,
                                  ctypes.byref(dev) if keep_on_device else None, ctypes.byref(ms)))
    if keep_on_device:
        return dev.value, n, H, W, ms.value
    masks.last_ms = ms.value
    return outs


def device_free(ptr, device=None):
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    _cabi.check(_cabi.lib().bp_device_free(int(device), ctypes.c_void_p(ptr)))

"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the reference's mask producer.

Follows /root/reference/plant3dvision/proc2d.py:69-115 (``linear``), :118-169 (``excess_green``), :172-220
(``dilation``) and /root/reference/plant3dvision/tasks/proc2d.py:224-237 (``Masks.f``: filter, ``> threshold``,
dilation, ``255 * img`` as uint8) statement by statement with numpy.

Third-party arithmetic: the reference calls scikit-image (unpinned in its requirements; ABSENT from this image)
for ``skimage.exposure.rescale_intensity(img, out_range=(0., 1.))`` and
``skimage.morphology.binary_dilation(img, footprint=disk(n, decomposition='sequence'))``.  Restated here from
their published algorithms:
  * rescale_intensity with the default ``in_range='image'`` on a float64 image: ``imin, imax = image.min(),
    image.max()``; ``image = clip(image, imin, imax)``; if ``imin != imax``: ``(image - imin) / (imax - imin)``
    then ``* (omax - omin) + omin``; else ``clip(image, omin, omax)``;
  * binary_dilation with a footprint sequence: ``scipy.ndimage.binary_dilation(., structure=fp,
    iterations=k, border_value=0)`` for every ``(fp, k)`` of the sequence (scipy IS in this image and is called).
  * ``disk(n, decomposition='sequence')`` reads a table shipped inside scikit-image (``disk_decompositions.npy``)
    that cannot be restated: ``disk_sequence`` returns it from skimage when importable, the exact cross for n = 1
    (documented: radius 1 is not decomposed), and raises otherwise.  Callers hand the sequence over explicitly.
PARITY UNPINNED for the two skimage functions (no skimage here to run); everything else is numpy / scipy itself.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import numpy as np
from scipy import ndimage as ndi

EPS = 1e-9  # proc2d.py:22


def rescale_intensity_01(image):
    imin, imax = float(np.min(image)), float(np.max(image))
    omin, omax = 0.0, 1.0
    image = np.clip(image, imin, imax)
    if imin != imax:
        image = (image - imin) / (imax - imin)
        return np.asarray(image * (omax - omin) + omin, dtype=np.float64)
    return np.clip(image, omin, omax).astype(np.float64)


def linear(img, coefs):
    """proc2d.py:111-115."""
    if not img.dtype == "float":
        img = np.asarray(img, dtype=float)
    img = rescale_intensity_01(img)
    return coefs[0] * img[:, :, 0] + coefs[1] * img[:, :, 1] + coefs[2] * img[:, :, 2]


def excess_green(img):
    """proc2d.py:161-169."""
    if not img.dtype == "float":
        img = np.asarray(img, dtype=float)
    img = rescale_intensity_01(img)
    s = img.sum(axis=2) + EPS
    r = img[:, :, 0] / s
    g = img[:, :, 1] / s
    b = img[:, :, 2] / s
    return (2 * g - r - b)


def disk_sequence(n):
    """``skimage.morphology.disk(n, decomposition='sequence')`` as ((footprint, repeats), ...)."""
    try:
        from skimage.morphology import disk
        return tuple((np.asarray(fp), int(k)) for fp, k in disk(n, decomposition="sequence"))
    except ImportError:
        if n == 1:
            return ((np.array([[0, 1, 0], [1, 1, 1], [0, 1, 0]], dtype=np.uint8), 1),)
        raise


def binary_dilation_sequence(img, sequence):
    """skimage.morphology.binary_dilation(img, footprint=<sequence>) (called at proc2d.py:219)."""
    out = np.asarray(img) != 0
    for fp, k in sequence:
        if k > 0:
            out = ndi.binary_dilation(out, structure=np.asarray(fp), iterations=int(k), border_value=0)
    return out


def masks_filter(img, type="linear", parameters=(0, 1, 0), threshold=0.3, dilation_sequence=()):
    """tasks/proc2d.py:213-237 for one image -> uint8 0 / 255."""
    if type == "linear":
        img = linear(img, list(parameters))
    elif type == "excess_green":
        img = excess_green(img)
    else:
        raise Exception(f"Unknown masking type '{type}'!")
    img = img > threshold
    if len(dilation_sequence) > 0:
        img = binary_dilation_sequence(img, dilation_sequence)
    return np.array(255 * img, dtype=np.uint8)

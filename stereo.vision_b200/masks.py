"""Masks of the path (functions.py:29-35).

The reference reads five PNGs relative to the current directory at import time and derives
carmask = car_front_mask where view_range != 0 (functions.py:35).  Only `carmask` feeds the hot path
(maskDisparity); `road_threshold_mask` is used by the road-mask cleaning that follows it.  Like the
reference, masks/*.png in the current directory win when present; otherwise the arrays derived from
the reference's shipped masks (data/masks_544x1024.npz, made by oracle/make_golden.py) are used,
nearest-neighbour scaled for frames that are an integer multiple of 544x1024.
"""
import os

import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "masks_544x1024.npz")
_cache = {}


def _shipped():
    if "npz" not in _cache:
        with np.load(_DATA) as z:
            _cache["npz"] = {k: z[k] for k in z.files}
    return _cache["npz"]


def _from_cwd():
    """carmask from masks/*.png in the CWD, exactly as functions.py:31-35 builds it."""
    a, b = "masks/car_front_mask.png", "masks/view_range.png"
    if not (os.path.isfile(a) and os.path.isfile(b)):
        return None
    try:
        import cv2
    except ImportError:
        return None
    cf = cv2.imread(a, cv2.IMREAD_GRAYSCALE)
    vr = cv2.imread(b, cv2.IMREAD_GRAYSCALE)
    if cf is None or vr is None:
        return None
    return np.where(vr != 0, cf, 0).astype(np.uint8)


def _fit(mask, shape):
    H, W = shape
    if mask.shape == (H, W):
        return mask
    if H % mask.shape[0] == 0 and W % mask.shape[1] == 0 and H // mask.shape[0] == W // mask.shape[1]:
        s = H // mask.shape[0]
        return np.repeat(np.repeat(mask, s, axis=0), s, axis=1)
    raise ValueError("no mask for frames of shape %s: pass carmask= explicitly" % (shape,))


def default_carmask(shape=(544, 1024), prefer_cwd=True):
    m = _from_cwd() if prefer_cwd else None
    if m is None:
        m = _shipped()["carmask"]
    return np.ascontiguousarray(_fit(m, shape))


def default_road_threshold_mask(shape=(544, 1024)):
    return np.ascontiguousarray(_fit(_shipped()["road_threshold_mask"], shape))

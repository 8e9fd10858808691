"""TEST INFRASTRUCTURE -- not part of the product.

ctypes face of oracle/sgbm_oracle.c: the CPU restatement of the reference's disparity stage
(functions.py:26, 104-128 = cv2.StereoSGBM_create(0, 128, 21).compute, cv2.filterSpeckles, scaling).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libsvb_oracle.so")
        if not os.path.isfile(path):
            import build_oracle
            build_oracle.build()
        _LIB = ctypes.CDLL(path)
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def sgbm_compute(grayL, grayR, num_disparities=128, block=21, median=True):
    """StereoSGBM_create(0, num_disparities, block).compute(grayL, grayR): int16 [H,W], disparity x 16,
    -16 where invalid (functions.py:26,108)."""
    L = np.ascontiguousarray(grayL, np.uint8)
    R = np.ascontiguousarray(grayR, np.uint8)
    assert L.ndim == 2 and L.shape == R.shape
    H, W = L.shape
    out = np.empty((H, W), np.int16)
    rc = lib().svo_sgbm(_p(L), _p(R), H, W, int(num_disparities), int(block), 2, 5, 1, 0, 15, int(bool(median)), _p(out))
    if rc != 0:
        raise ValueError("image narrower than numDisparities + half a block")
    return out


def filter_speckles(img16, new_val=0, max_size=4000, max_diff=123):
    """cv2.filterSpeckles on int16, in place like cv2 (functions.py:111); returns the array."""
    assert img16.dtype == np.int16 and img16.flags.c_contiguous
    H, W = img16.shape
    lib().svo_filter_speckles(_p(img16), H, W, int(new_val), int(max_size), int(max_diff))
    return img16


def scale_disparity(disp16, max_disparity=128, crop_disparity=False):
    """functions.py:117-127: threshold to zero, /16. -> uint8 (truncation), optional crop, x (256/max) -> uint8."""
    d = np.where(disp16 > 0, disp16, 0).astype(np.int16)
    s = (d / 16.).astype(np.uint8)
    if crop_disparity:
        s = s[0:390, 135:s.shape[1]]
    return (s * (256. / max_disparity)).astype(np.uint8)


def disparity(grayL, grayR, max_disparity=128, crop_disparity=False, block=21):
    """functions.disparity (functions.py:104-128) with the module's matcher (functions.py:26)."""
    d16 = sgbm_compute(grayL, grayR, 128, block)
    filter_speckles(d16, 0, 4000, max_disparity - 5)
    return scale_disparity(d16, max_disparity, crop_disparity)

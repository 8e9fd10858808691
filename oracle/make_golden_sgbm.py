#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- writes tests/golden/sgbm.npz from the UNMODIFIED reference.

    python oracle/make_golden_sgbm.py

For every input pair of tests/helpers.sgbm_inputs() it calls the reference's own functions.disparity
(functions.py:104-128, with the module's StereoSGBM_create(0, 128, 21), functions.py:26) and, for the
intermediate products, the same cv2 calls one by one.  Recorded: sha-256 of the inputs (so drift of the
synthetic generator is caught), sha-256 of the raw int16 matcher output, of the speckle-filtered int16 image
and of the final uint8 disparity, the final disparity itself for the cases up to 120x400, and counts.
Runs only in the build container (/root/reference and cv2 are absent on the GPU box).
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_harness as rh          # noqa: E402
from helpers import sgbm_inputs   # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    import cv2
    f = rh.load_reference()
    flat, names = {}, []
    for name, L, R, crop in sgbm_inputs():
        final = f.disparity(L, R, 128, crop)
        raw = f.stereoProcessor.compute(L, R)
        spk = raw.copy()
        cv2.filterSpeckles(spk, 0, 4000, 128 - 5)
        p = name + "_"
        flat[p + "in_sha"] = sha(L) + sha(R)
        flat[p + "raw_sha"], flat[p + "speckle_sha"], flat[p + "final_sha"] = sha(raw), sha(spk), sha(final)
        small = raw.copy()                       # a speckle size that leaves something standing on the small inputs
        cv2.filterSpeckles(small, 0, 60, 40)
        flat[p + "speckle60_sha"] = sha(small)
        flat[p + "shape"] = np.array(final.shape)
        flat[p + "valid"] = int((raw >= 0).sum())
        flat[p + "despeckled"] = int(((raw != 0) & (spk == 0)).sum())
        flat[p + "nonzero"] = int((final != 0).sum())
        if L.size <= 120 * 400:
            flat[p + "final"] = final
            flat[p + "raw"] = raw
        names.append(name)
        print("  %-10s %s valid %d, despeckled %d, nonzero %d, max %d" % (name, L.shape, flat[p + "valid"], flat[p + "despeckled"],
                                                                        flat[p + "nonzero"], int(final.max())))
    flat["names"] = np.array(names)
    flat["cv2_version"] = cv2.__version__
    path = os.path.join(ROOT, "tests", "golden", "sgbm.npz")
    np.savez_compressed(path, **flat)
    print("wrote sgbm.npz (%.1f KB)" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()

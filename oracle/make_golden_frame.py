#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- writes tests/golden/frames.npz from the UNMODIFIED reference orchestrator.

    python oracle/make_golden_frame.py

Executes /root/reference/stereovision.py (source untouched, its `import functions as f` bound to the reference's
own functions.py) on three chained synthetic BGR stereo frames with `random.seed(synth.ransac_seed(seq, t))`
before each call, options as loop.py:12-26 with 'loop': False (no window).  The helpers the orchestrator calls
are wrapped by recorders -- module attributes of the loaded `functions` module are replaced by pass-through
closures, no file of the reference is modified -- so that every stage's product is captured: the matcher's
disparity, the filled disparity that is carried to the next frame, the plane, the point counts, the road image,
the cleaned image with its points, the hull centre.  cv2 is OpenCV 4 here, so `findContours` is presented with the
OpenCV-3 return signature the reference was written for (ref_harness._Cv2Compat), as for the sanitise fixture.
"""
import hashlib
import os
import random
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_harness as rh          # noqa: E402
import importlib                  # noqa: E402
synth = importlib.import_module("stereo.vision_b200.synth")

SEQ, FRAMES = 2, 3


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load_orchestrator(f):
    sys.modules["functions"] = f
    path = os.path.join(rh.REF_DIR, "stereovision.py")
    mod = types.ModuleType("reference_stereovision")
    mod.__file__ = path
    with open(path) as fh:
        exec(compile(fh.read(), path, "exec"), mod.__dict__)
    return mod


def main():
    f = rh.load_reference()
    f.cv2 = rh._Cv2Compat(f.cv2)
    rec = {}

    def record(name):
        orig = getattr(f, name)

        def wrapper(*a, **k):
            out = orig(*a, **k)
            rec.setdefault(name, []).append(out)
            return out
        setattr(f, name, wrapper)
    for name in ("disparity", "disparityCleaning", "RANSAC", "computePlanarThreshold", "filterPointsByHistogram",
                 "generatePointsAsImage", "sanitiseRoadImage", "getCenterPoint", "getNormalVectorLine"):
        record(name)
    sv = load_orchestrator(f)
    opts = {'crop_disparity': False, 'pause_playback': False, 'max_disparity': 128, 'ransac_trials': 600, 'road_color_thresh': 10,
            'loop': False, 'point_threshold': 0.05, 'image_tiles': True, 'img_size': (544, 1024), 'threshold_option': 'previous',
            'record_video': False, 'record_stats': False, 'video_filename': 'previous.avi'}
    flat = {"seq": SEQ, "n_frames": FRAMES}
    prev = None
    for t in range(FRAMES):
        rec.clear()
        imgL, imgR = synth.make_stereo_bgr(SEQ, t)
        random.seed(synth.ransac_seed(SEQ, t))
        cwd = os.getcwd()
        os.chdir(rh.REF_DIR)
        try:
            _, prev, normal = sv.performStereoVision(imgL.copy(), imgR.copy(), prev, opts)
        finally:
            os.chdir(cwd)
        p = "f%d_" % t
        flat[p + "in_sha"] = sha(imgL) + sha(imgR)
        flat[p + "disparity_sha"] = sha(rec["disparity"][0])
        flat[p + "filled_sha"] = sha(prev)
        flat[p + "disparity_nonzero"] = int((rec["disparity"][0] != 0).sum())
        flat[p + "has_plane"] = normal is not None
        if normal is not None:
            flat[p + "abc"] = np.asarray(normal, np.float64).reshape(3)
            flat[p + "near_n"] = len(rec["computePlanarThreshold"][0])
            flat[p + "road_n"] = len(rec["filterPointsByHistogram"][0])
        road = rec["generatePointsAsImage"][0]
        flat[p + "road_sha"], flat[p + "road_px"] = sha(road), int((road != 0).sum())
        if "sanitiseRoadImage" in rec:
            cleaned, pts = rec["sanitiseRoadImage"][0]
            flat[p + "cleaned_sha"], flat[p + "cleaned_pts"] = sha(cleaned), int(np.asarray(pts).reshape(-1, 2).shape[0] if np.asarray(pts).size else 0)
        if "getCenterPoint" in rec:
            flat[p + "centre"] = np.array(rec["getCenterPoint"][0])
        if "getNormalVectorLine" in rec:
            flat[p + "normal_line"] = np.array(rec["getNormalVectorLine"][0])
        print("  frame %d: disparity nonzero %d, plane %s, near %s, road %s, road px %d, cleaned pts %s, centre %s, line %s" % (
            t, flat[p + "disparity_nonzero"], None if normal is None else np.round(flat[p + "abc"], 4), flat.get(p + "near_n"),
            flat.get(p + "road_n"), flat[p + "road_px"], flat.get(p + "cleaned_pts"), flat.get(p + "centre"), flat.get(p + "normal_line")))
    path = os.path.join(ROOT, "tests", "golden", "frames.npz")
    np.savez_compressed(path, **flat)
    print("wrote frames.npz (%.1f KB)" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()

"""TEST INFRASTRUCTURE -- not part of the product.

Runs the *unmodified* reference helpers (/root/reference/functions.py) on in-memory arrays, in the
order the orchestrator calls them (stereovision.py:58-140), and captures every intermediate.  The
reference tree only exists in the build container, so this module is used by exactly one thing:
`oracle/make_golden.py`, which turns its outputs into the fixtures under tests/golden/.  Nothing
under tests/ marked gpu, smoke() or bench.py imports it.

Two knobs of the reference are literals in its source: the lattice step (functions.py:185-186, "2")
and the RANSAC sample size (functions.py:286, "600").  To obtain reference outputs for the other
BASELINE configurations the loader can substitute those two literals in the source text it executes
(never on disk); with the defaults the source is executed untouched.
"""
import os
import random
import sys
import types
import warnings

import numpy as np

REF_DIR = "/root/reference"


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_DIR, "functions.py"))


def load_reference(step: int = 2, n_samples: int = 600):
    """Execute functions.py (CWD = reference dir, its masks are CWD-relative, functions.py:29-33)."""
    with open(os.path.join(REF_DIR, "functions.py")) as fh:
        src = fh.read()
    if step != 2:
        a = "for y in range(0,height-1, 2)"
        b = "for x in range(0,width-1, 2)"
        assert a in src and b in src
        src = src.replace(a, "for y in range(0,height-1, %d)" % step)
        src = src.replace(b, "for x in range(0,width-1, %d)" % step)
    if n_samples != 600:
        a = "random.sample(points, 600)"
        assert a in src
        src = src.replace(a, "random.sample(points, %d)" % n_samples)
    mod = types.ModuleType("reference_functions")
    mod.__file__ = os.path.join(REF_DIR, "functions.py")
    cwd = os.getcwd()
    os.chdir(REF_DIR)
    try:
        exec(compile(src, mod.__file__, "exec"), mod.__dict__)
    finally:
        os.chdir(cwd)
    return mod


def install_masks(f, carmask=None, black=None):
    """Swap the module-global masks (functions.py:32,35), e.g. for the 2048x1088 stress frames."""
    if carmask is not None:
        f.carmask = np.ascontiguousarray(carmask)
    if black is not None:
        f.blackImg = np.ascontiguousarray(black)
        # capDisparity (functions.py:166) masks with disparity_range; its result is discarded but the
        # shapes must agree for the call to succeed
        f.disparity_range = np.full(black.shape, 255, dtype=np.uint8)


def install_camera(f, cam):
    f.camera_focal_length_px = cam.focal_px
    f.stereo_camera_baseline_m = cam.baseline_m
    f.image_centre_h = cam.centre_h
    f.image_centre_w = cam.centre_w


def replay_indices(points_xyz: np.ndarray, trials: int, n_samples: int, seed: int):
    """Replay the reference's draw sequence on index ranges.

    Per trial: random.sample(points, S) (functions.py:286) then rounds of three
    random.sample(points, 1) (functions.py:252-254) until the cross product is not exactly zero
    (functions.py:251,256-259).  Sampling a list and sampling range(len(list)) consume the generator
    identically, so this yields the indices the reference picked under the same seed.
    Returns (sample_idx [T,S] int32, triple_idx [T,3] int32, rounds [T] int32); all -1 when M < S.
    """
    m = len(points_xyz)
    rng = random.Random()
    rng.seed(seed)
    samp = -np.ones((trials, n_samples), dtype=np.int32)
    trip = -np.ones((trials, 3), dtype=np.int32)
    rounds = np.zeros(trials, dtype=np.int32)
    if m < n_samples:
        return samp, trip, rounds
    pop = range(m)
    for t in range(trials):
        samp[t] = rng.sample(pop, n_samples)
        while True:
            i1 = rng.sample(pop, 1)[0]
            i2 = rng.sample(pop, 1)[0]
            i3 = rng.sample(pop, 1)[0]
            rounds[t] += 1
            cp = np.cross(points_xyz[i1] - points_xyz[i2], points_xyz[i2] - points_xyz[i3])
            if not (cp[0] == 0 and cp[1] == 0 and cp[2] == 0):
                break
        trip[t] = (i1, i2, i3)
    return samp, trip, rounds


def run_frame(f, disparity, bgr, prev_disp, seed, trials=600, point_threshold=0.05,
              road_color_thresh=10):
    """The helper chain of stereovision.py:58-140 on one frame.  Returns a dict of intermediates."""
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        disp = f.disparityCleaning(disparity.copy(), 'previous', prev_disp)     # stereovision.py:58
        out["filled"] = np.array(disp, copy=True)
        masked = f.maskDisparity(disp)                                            # :74
        capped = f.capDisparity(disp)                                             # :76
        assert capped is disp
        out["masked_disp"] = np.array(masked, copy=True)
        points = f.projectDisparityTo3d(capped, 128, bgr)                         # :84
        maskpoints = f.projectDisparityTo3d(masked, 128)                          # :85
        out["points_xyz"] = np.array([p[:3] for p in points], dtype=np.float64).reshape(-1, 3)
        out["points_rgb"] = np.array([p[3:6] for p in points], dtype=np.uint8).reshape(-1, 3)
        out["mask_xyz"] = np.array(maskpoints, dtype=np.float64).reshape(-1, 3)

        # RANSAC (:94) under the frame's seed, recording each trial's plane and error
        log = []
        orig_fit = f.planarFitting

        def spy(T, pts):
            abc, nrm, dist = orig_fit(T, pts)
            log.append((np.array(abc, dtype=np.float64).reshape(3), float(np.mean(dist))))
            return abc, nrm, dist
        f.planarFitting = spy
        try:
            random.seed(seed)
            normal, abc = f.RANSAC(maskpoints, trials)
        finally:
            f.planarFitting = orig_fit
        out["trial_log_n"] = len(log)
        if abc is None:
            out["abc"] = None
            out["best_error"] = np.inf
            out["near_n"] = 0
            out["road_n"] = 0
            out["hist"] = {}
            out["plane_pts_int"] = np.zeros((0, 2), dtype=np.int32)
            out["road_mask"] = np.array(f.blackImg, copy=True)
            return out
        out["abc"] = np.array(abc, dtype=np.float64).reshape(3)
        errs = np.array([e for _, e in log])
        out["best_error"] = float(errs.min())
        diffs = f.calculatePointErrors(abc, points)                               # :97
        out["diffs"] = np.array(diffs, dtype=np.float64).reshape(-1)
        near = f.computePlanarThreshold(points, diffs, point_threshold)           # :100
        out["near_n"] = len(near)
        hist = f.calculateColourHistogram(near)                                   # :103
        out["hist"] = dict(hist)
        road = f.filterPointsByHistogram(near, hist, road_color_thresh)           # :106
        out["road_n"] = len(road)
        pp = f.project3DPointsTo2DImagePoints(road)                               # :111
        pp = np.array(pp, np.int32).reshape((-1, 1, 2))                           # :112-113
        out["plane_pts_int"] = pp.reshape(-1, 2).copy()
        out["road_mask"] = f.generatePointsAsImage(pp)                            # :140
    return out



class _Cv2Compat:
    """cv2 as the reference expects it: functions.py:379 unpacks three values from findContours
    (the OpenCV 3 signature); OpenCV 4 returns (contours, hierarchy).  Everything else passes through."""

    def __init__(self, cv2):
        self._cv2 = cv2

    def __getattr__(self, name):
        return getattr(self._cv2, name)

    def findContours(self, image, mode, method, **kw):
        res = self._cv2.findContours(image, mode, method, **kw)
        if len(res) == 2:
            return image, res[0], res[1]
        return res


def run_sanitise(f, road_image, road_threshold_mask=None):
    """stereovision.py:156: the unmodified sanitiseRoadImage (functions.py:346-366) on one road image.
    Returns (cleaned uint8 image, pts int32 [N,2] as (x,y), calls made by ParticleCleansing's fill)."""
    if not isinstance(f.cv2, _Cv2Compat):
        f.cv2 = _Cv2Compat(f.cv2)
    if road_threshold_mask is not None:
        f.road_threshold_mask = np.ascontiguousarray(road_threshold_mask)
    H, W = road_image.shape
    img, pts = f.sanitiseRoadImage(np.array(road_image, copy=True), (H, W))
    pts = np.asarray(pts, dtype=np.int32).reshape(-1, 2) if pts.size else np.zeros((0, 2), np.int32)
    return np.asarray(img), pts


def run_preprocess(f, imgL, imgR):
    """stereovision.py:44,46: the unmodified preProcessImages + greyscale (functions.py:80-96)."""
    gl, gr = f.preProcessImages(np.array(imgL, copy=True), np.array(imgR, copy=True))
    yl, yr = f.greyscale(gl, gr)
    return np.asarray(gl), np.asarray(gr), np.asarray(yl), np.asarray(yr)


def run_fill_alt(f, disparity):
    """stereovision.py:58 with threshold_option 'mean': disparityCleaning -> fillAltDisparity (functions.py:136-137,
    149-162) on a copy (the reference writes into its argument)."""
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")          # mean of an empty slice
        return np.asarray(f.disparityCleaning(np.array(disparity, copy=True), 'mean', None))


def run_normal_line(f, base_point, abc, disparity):
    """functions.py:396-416 on a (3,1) plane as RANSAC returns it; (x, y) or the exception's class name."""
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return tuple(int(v) for v in f.getNormalVectorLine(base_point, np.asarray(abc, dtype=np.float64).reshape(3, 1), disparity))
    except Exception as e:                       # int(nan) -> ValueError, int(inf) -> OverflowError
        return type(e).__name__


def run_obstacles(f, cleaned, pts):
    """stereovision.py:170-200: the orchestrator's own cv2 calls (these lines are not helpers of functions.py),
    then the reference's drawRoadLine / getCenterPoint.  Returns (hull (n,2), hull mask, obstacle mask, centre)."""
    cv2 = f.cv2
    pts = np.asarray(pts, dtype=np.int32).reshape(-1, 1, 2)
    roadHull = cv2.convexHull(pts)                                                   # :172
    hullMask = cv2.drawContours(cleaned.copy(), [roadHull], 0, 255, -100)           # :174,176
    obstacleImage = cv2.bitwise_not(cleaned.copy())                                  # :175,178
    obstacleImage = cv2.bitwise_and(obstacleImage, obstacleImage, mask=hullMask)     # :180
    _, hull = f.drawRoadLine(np.zeros(cleaned.shape + (3,), np.uint8), pts)          # :198 (functions.py:390-394)
    centre = f.getCenterPoint(hull)                                                  # :200 (functions.py:418-421)
    assert np.array_equal(hull, roadHull)
    return roadHull.reshape(-1, 2), hullMask, obstacleImage, tuple(int(v) for v in centre)

def hist_to_bins(hist: dict) -> np.ndarray:
    """Reference histogram dict (keys = str(round(hue,3))) -> 1001 counts indexed by rint(hue*1000)."""
    bins = np.zeros(1001, dtype=np.int64)
    for k, v in hist.items():
        b = int(round(float(k) * 1000.0))
        assert 0 <= b <= 1000 and bins[b] == 0, k
        bins[b] = v
    return bins


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import importlib
    svb = importlib.import_module("stereo.vision_b200.synth")
    f = load_reference()
    d, bgr = svb.make_frame(0, 0)
    import time
    t0 = time.perf_counter()
    o = run_frame(f, d, bgr, None, svb.ransac_seed(0, 0))
    print("ref frame %.2fs" % (time.perf_counter() - t0), {k: (v.shape if hasattr(v, "shape") else v) for k, v in o.items() if k != "hist"})

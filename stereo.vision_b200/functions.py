"""Drop-in for the hot-path helpers of the reference's functions.py (same names, positional
signatures and error behaviour), executed by the sm_100a kernels behind libsvb200.so.

    import stereo.vision_b200.functions as f        # instead of `import functions as f`

Covered: the calls of stereovision.py:58-140 -- disparityCleaning / fillDisparity, maskDisparity,
capDisparity, projectDisparityTo3d, RANSAC, calculatePointErrors, computePlanarThreshold,
calculateColourHistogram, filterPointsByHistogram, project3DPointsTo2DImagePoints,
generatePointsAsImage -- plus the module constants they read, the first consumer of the road image,
sanitiseRoadImage (stereovision.py:156), and the front end gammaChange / preProcessImages / greyscale
(stereovision.py:44,46).  Clouds are returned as device handles
(handles.py) that behave like the reference's lists for everything the orchestrator does with them.
There is no CPU fallback: without a CUDA device every helper raises.

Two literals of the reference are module settings here: `lattice_step` (functions.py:185-186,
2) and `ransac_samples` (functions.py:286, 600).
"""
import ctypes as C
import random

import numpy as np
import torch

from . import _native as N
from .camera import DEFAULT_CAMERA, Camera
from .handles import DeviceColumn, HueHistogram, ImagePoints, PointCloud, bins_from_dict
from .masks import default_carmask, default_road_threshold_mask

# ---- constants, functions.py:15-24
camera_focal_length_px = DEFAULT_CAMERA.focal_px
camera_focal_length_m = 4.8 / 1000
stereo_camera_baseline_m = DEFAULT_CAMERA.baseline_m
image_centre_h = DEFAULT_CAMERA.centre_h
image_centre_w = DEFAULT_CAMERA.centre_w
max_disparity = 128
# ---- masks, functions.py:29-35.  carmask: masks/car_front_mask.png + masks/view_range.png of the CWD when present
# (as the reference reads them), else the packaged arrays; road_threshold_mask and blackImg: the packaged arrays
# (assign to the module attributes to use others)
carmask = default_carmask((544, 1024))
blackImg = np.zeros((544, 1024), dtype=np.uint8)
road_threshold_mask = default_road_threshold_mask((544, 1024))
# ---- literals of the reference exposed as settings
lattice_step = 2
ransac_samples = 600
ransac_replay = "device"      # where Python's `random` stream is replayed: "device" (svb_ransac_replay_draws) or "host"

_state = {}


def _camera() -> Camera:
    return Camera(camera_focal_length_px, stereo_camera_baseline_m, image_centre_h, image_centre_w)


def _dev():
    N.require_cuda()
    return torch.device("cuda", torch.cuda.current_device())


def _to_dev_u8(a):
    if torch.is_tensor(a):
        return a.to(_dev()).contiguous(), True
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint8)).to(_dev()), False


def _carmask_dev():
    """`carmask` on the current device.  The cache key holds a digest of the uploaded bytes, so an in-place edit of the
    module's array is seen, and there is one copy per device."""
    import hashlib
    dev = _dev()
    a = np.ascontiguousarray(carmask, dtype=np.uint8)
    key = (str(dev), a.shape, hashlib.blake2b(a.tobytes(), digest_size=16).digest())
    if _state.get("cm_key") != key:
        _state["cm"] = torch.from_numpy(a).to(dev)
        _state["cm_key"] = key
    return _state["cm"]


def _same_size(what, a_shape, b_shape):
    """cv2 raises on operands of different sizes (functions.py:142-146,171), and the orchestrator relies on it: stage 2
    then falls back to the black image (stereovision.py:54-64)."""
    if tuple(a_shape[:2]) != tuple(b_shape[:2]):
        raise ValueError("%s: operand sizes differ, %s and %s" % (what, tuple(a_shape[:2]), tuple(b_shape[:2])))


def _back(t, was_tensor):
    return t if was_tensor else t.cpu().numpy()


# ---------------------------------------------------------------- a1-a3
def disparityCleaning(disparity, option, prev_disp=None):
    """functions.py:130-138."""
    if option == 'previous':
        if prev_disp is not None:
            disparity = fillDisparity(disparity, prev_disp)
    elif option == 'mean':
        disparity = fillAltDisparity(disparity)
    return disparity


def fillDisparity(disparity, previousDisparity):
    """functions.py:140-147."""
    if previousDisparity is None:
        return disparity
    d, was = _to_dev_u8(disparity)
    p, _ = _to_dev_u8(previousDisparity)
    if d.dim() != 2:
        raise ValueError("fillDisparity expects a single-channel disparity")
    _same_size("fillDisparity", d.shape, p.shape)
    H, W = d.shape
    out = torch.empty_like(d)
    N.check(N.load().svb_fill_sequence(N.ptr(d), N.ptr(p), N.ptr(out), 1, H, W, N.stream_ptr()), "svb_fill_sequence")
    N.count(1)
    return _back(out, was)


def capDisparity(disparity):
    """functions.py:164-167: the masked copy is discarded; the argument is returned unchanged."""
    return disparity


def maskDisparity(disparity):
    """functions.py:169-172."""
    d, was = _to_dev_u8(disparity)
    if d.dim() != 2:
        raise ValueError("maskDisparity expects a single-channel disparity")
    _same_size("maskDisparity (carmask)", d.shape, np.shape(carmask))
    H, W = d.shape
    out = torch.empty_like(d)
    N.check(N.load().svb_mask_disparity(N.ptr(d), N.ptr(_carmask_dev()), N.ptr(out), 1, H, W, N.stream_ptr()),
            "svb_mask_disparity")
    N.count(1)
    return _back(out, was)


# ---------------------------------------------------------------- a4
def projectDisparityTo3d(disparity, max_disparity, rgb=[]):
    """functions.py:178-198.  Returns a PointCloud (len / iteration / indexing / np.array like the list)."""
    lib = N.load()
    d, _ = _to_dev_u8(disparity)
    H, W = d.shape[:2]
    has_rgb = len(rgb) > 0
    c = _to_dev_u8(rgb)[0] if has_rgb else None
    if has_rgb and (c.dim() != 3 or c.shape[2] != 3 or tuple(c.shape[:2]) != (H, W)):
        raise IndexError("projectDisparityTo3d: rgb is %s, the disparity %s" % (tuple(c.shape), (H, W)))    # rgb[y,x,2], :195
    step = int(lattice_step)
    L = lib.svb_lattice_rows(H, step) * lib.svb_lattice_cols(W, step)
    dev = d.device
    cloud = torch.empty((max(L, 1),), dtype=torch.int64, device=dev)
    counts = torch.zeros((1, 4), dtype=torch.int32, device=dev)
    scratch = torch.empty((max(16, lib.svb_project_scratch_bytes(1, H, W, step)),), dtype=torch.uint8, device=dev)
    cam = _camera()
    cc = N.ccam(cam)
    N.check(lib.svb_project_compact(N.ptr(d), None, N.ptr(c), 1, H, W, step, C.byref(cc), N.ptr(cloud), None,
                                    N.ptr(counts), N.ptr(scratch), scratch.numel(), N.stream_ptr()), "svb_project_compact")
    N.count(2)
    n = int(counts[0, 0].item())
    return PointCloud(cloud, n, cam, has_rgb)


def project3DPointsTo2DImagePoints(points):
    """functions.py:201-209."""
    _need_cloud(points)
    xy = torch.empty((len(points), 2), dtype=torch.float64, device=points.records.device)
    cc = N.ccam(points.camera)
    N.check(N.load().svb_reproject(N.ptr(points.records), 2, len(points), C.byref(cc), N.ptr(xy), None, N.stream_ptr()),
            "svb_reproject")
    N.count(1 if len(points) else 0)
    return ImagePoints(xy)


def _need_cloud(points):
    if not isinstance(points, PointCloud):
        raise TypeError("expected the PointCloud returned by projectDisparityTo3d, got %s" % type(points).__name__)


# ---------------------------------------------------------------- a10-a12
def calculateColourHistogram(points):
    """functions.py:215-226."""
    _need_cloud(points)
    hist = torch.zeros((N.HIST_STRIDE,), dtype=torch.int32, device=points.records.device)
    N.check(N.load().svb_hue_histogram(N.ptr(points.records), len(points), N.ptr(hist), N.stream_ptr()), "svb_hue_histogram")
    N.count(1 if len(points) else 0)
    return HueHistogram(hist)


def _compact(points, mode, dist, hist, threshold):
    lib, n = N.load(), len(points)
    dev = points.records.device
    out = torch.empty((max(n, 1),), dtype=torch.int64, device=dev)
    cnt = torch.zeros((1,), dtype=torch.int32, device=dev)
    scratch = torch.empty((max(16, lib.svb_compact_scratch_bytes(n)),), dtype=torch.uint8, device=dev)
    N.check(lib.svb_compact_cloud(N.ptr(points.records), n, mode, N.ptr(dist), N.ptr(hist), float(threshold), N.ptr(out),
                                  N.ptr(cnt), N.ptr(scratch), scratch.numel(), N.stream_ptr()), "svb_compact_cloud")
    N.count(3 if n else 0)
    return PointCloud(out, int(cnt.item()), points.camera, points.has_rgb)


def filterPointsByHistogram(points, histogram, threshold=100):
    """functions.py:228-230: keep points whose hue key count is > threshold, order kept."""
    _need_cloud(points)
    return _compact(points, 1, None, bins_from_dict(histogram, points.records.device), threshold)


# ---------------------------------------------------------------- a8-a9
def calculatePointErrors(abc, points):
    """functions.py:300-312.  abc None raises TypeError like `abc[0]` does in the reference (:307)."""
    _need_cloud(points)
    if abc is None:
        raise TypeError("'NoneType' object is not subscriptable")
    h = np.ascontiguousarray(np.asarray(abc, dtype=np.float64).reshape(-1)[:3])
    dist = torch.empty((len(points),), dtype=torch.float64, device=points.records.device)
    cc = N.ccam(points.camera)
    N.check(N.load().svb_point_errors(N.ptr(points.records), 2, len(points), C.byref(cc),
                                      h.ctypes.data_as(C.c_void_p), N.ptr(dist), N.stream_ptr()), "svb_point_errors")
    N.count(1 if len(points) else 0)
    return DeviceColumn(dist)


def computePlanarThreshold(points, differences, threshold=0.01):
    """functions.py:314-323: keep points[i] where differences[i] < threshold, order kept."""
    _need_cloud(points)
    if isinstance(differences, DeviceColumn):
        dist = differences.t
    else:
        dist = torch.from_numpy(np.ascontiguousarray(np.asarray(differences, dtype=np.float64).reshape(-1))).to(points.records.device)
    if dist.numel() < len(points):
        raise IndexError("index %d is out of bounds for axis 0 with size %d" % (dist.numel(), dist.numel()))
    return _compact(points, 0, dist, None, threshold)


# ---------------------------------------------------------------- a5-a7
def _replay_draws(xyz, trials, S):
    """Consume Python's global `random` stream exactly as functions.py:286 and :252-254 do."""
    m = len(xyz)
    samp = np.empty((trials, S), dtype=np.int32)
    trip = np.empty((trials, 3), dtype=np.int32)
    pop = range(m)
    for t in range(trials):
        samp[t] = random.sample(pop, S)
        while True:
            i1, i2, i3 = random.sample(pop, 1)[0], random.sample(pop, 1)[0], random.sample(pop, 1)[0]
            u, v = xyz[i1] - xyz[i2], xyz[i2] - xyz[i3]
            if not (u[1] * v[2] - u[2] * v[1] == 0 and u[2] * v[0] - u[0] * v[2] == 0 and u[0] * v[1] - u[1] * v[0] == 0):
                break
        trip[t] = (i1, i2, i3)
    return samp, trip


def _device_draws(points, T, S, M):
    """The same draws as _replay_draws, made on the device from random.getstate() (svb_ransac_replay_draws);
    the global generator is left where the reference's RANSAC call would leave it (random.setstate)."""
    lib, dev = N.load(), points.records.device
    state = random.getstate()
    if state[0] != 3:
        raise N.SvbError("unexpected random.getstate() version %r" % (state[0],))
    mt = torch.from_numpy(np.array(state[1], dtype=np.uint32).view(np.int32)).to(dev).view(1, 625)
    si = torch.empty((1, T, S), dtype=torch.int32, device=dev)
    ti = torch.empty((1, T, 3), dtype=torch.int32, device=dev)
    nb = int(lib.svb_ransac_replay_scratch_bytes(1, S, M))
    scratch = torch.empty((nb,), dtype=torch.uint8, device=dev)
    counts = torch.tensor([[M]], dtype=torch.int32, device=dev)
    cc = N.ccam(points.camera)
    N.check(lib.svb_ransac_replay_draws(N.ptr(points.records), 2, 0, N.ptr(counts), 1, 0, 1, T, S, M, C.byref(cc), None,
                                        N.ptr(mt), 0, N.ptr(si), N.ptr(ti), N.ptr(scratch), nb, N.stream_ptr()),
            "svb_ransac_replay_draws")
    N.count(1)
    words = mt.cpu().numpy().view(np.uint32).reshape(-1)
    random.setstate((3, tuple(int(v) for v in words), state[2]))
    return si, ti


def RANSAC(points, trials):
    """functions.py:278-298.  Returns (normal, coefficients): the same (3,1) float64 array twice, or
    (None, None) when no trial produced a plane (e.g. fewer than `ransac_samples` points).  The draws are the
    reference's: Python's global `random` stream, replayed on the device (`ransac_replay = "host"` keeps the
    Python loop, its checker)."""
    _need_cloud(points)
    lib, S, T, M = N.load(), int(ransac_samples), int(trials), len(points)
    if M < S or T < 1:
        return (None, None)                   # random.sample raises in every trial, swallowed at :294
    dev = points.records.device
    if ransac_replay == "host":
        samp, trip = _replay_draws(points.xyz, T, S)
        si, ti = torch.from_numpy(samp).to(dev), torch.from_numpy(trip).to(dev)
    else:
        si, ti = _device_draws(points, T, S, M)
    counts = torch.tensor([[M]], dtype=torch.int32, device=dev)
    errors = torch.empty((1, T), dtype=torch.float64, device=dev)
    planes = torch.zeros((1, N.PLANE_STRIDE), dtype=torch.float64, device=dev)
    cc, st = N.ccam(points.camera), N.stream_ptr()
    N.check(lib.svb_ransac_score(N.ptr(points.records), 2, 0, N.ptr(counts), 1, 0, 1, T, S, M, C.byref(cc), N.ptr(si),
                                 N.ptr(ti), 0, 0, 0, N.ptr(errors), st), "svb_ransac_score")
    N.check(lib.svb_ransac_finalize(N.ptr(points.records), 2, 0, N.ptr(counts), 1, 0, 1, T, S, C.byref(cc), N.ptr(ti),
                                    0, 0, N.ptr(errors), 0.0, N.ptr(planes), st), "svb_ransac_finalize")
    N.count(2)
    row = planes[0].cpu().numpy()
    RANSAC.last = {"best_trial": int(row[6]), "best_error": float(row[5]), "errors": errors, "sample_idx": si, "triple_idx": ti}
    if row[7] == 0.0:
        return (None, None)
    abc = row[:3].reshape(3, 1).copy()
    return (abc, abc)


# ---------------------------------------------------------------- a14
def generatePointsAsImage(points):
    """functions.py:339-344: 255 at img[y][x] for each (x,y) of an int32 (R,1,2) array."""
    pts = np.ascontiguousarray(np.asarray(points, dtype=np.int32).reshape(-1, 2))
    H, W = blackImg.shape
    dev = _dev()
    img = torch.from_numpy(np.ascontiguousarray(blackImg)).to(dev)
    if len(pts):
        p = torch.from_numpy(pts).to(dev)
        oob = torch.zeros((1,), dtype=torch.int32, device=dev)
        val = (C.c_uint8 * 4)(255, 0, 0, 0)
        N.check(N.load().svb_scatter_points(N.ptr(p), len(pts), N.ptr(img), H, W, 1, val, N.ptr(oob), N.stream_ptr()),
                "svb_scatter_points")
        N.count(1)
        if int(oob.item()):
            raise IndexError("road point outside the %dx%d image" % (H, W))
    return img.cpu().numpy()


# ---------------------------------------------------------------- f1
def _threshold_bits(H, W):
    import hashlib
    if tuple(road_threshold_mask.shape) != (H, W):               # cv2.bitwise_and(.., mask=) raises on a size mismatch, :355
        raise ValueError("road_threshold_mask is %s, the image %s" % (road_threshold_mask.shape, (H, W)))
    a = np.ascontiguousarray(road_threshold_mask, dtype=np.uint8)
    key = (str(_dev()), a.shape, hashlib.blake2b(a.tobytes(), digest_size=16).digest())
    if _state.get("tb_key") != key:
        m = torch.from_numpy(a).to(_dev())
        bits = torch.empty((N.load().svb_bits_words(H, W),), dtype=torch.int32, device=m.device)
        N.check(N.load().svb_pack_mask(N.ptr(m), 1, H, W, N.ptr(bits), 0, N.stream_ptr()), "svb_pack_mask")
        N.count(1)
        _state["tb"], _state["tb_key"] = bits, key
    return _state["tb"]


def sanitiseRoadImage(img, size):
    """functions.py:346-366: closing, two erosions, road_threshold_mask, closing (9x9), then the (x,y)
    list of the non-zero pixels as an (N,2) matrix, rows outermost.  `size` = (height, width) as in the
    reference (opt['img_size']); the image must be the 0 / 255 mask generatePointsAsImage returns."""
    (height, width) = size
    t, was_tensor = _to_dev_u8(img)
    if t.dim() != 2:
        raise ValueError("sanitiseRoadImage expects a single-channel image")
    H, W = t.shape
    if height > H or width > W:
        raise IndexError("size %s exceeds the %dx%d image" % (size, H, W))      # img[i][j] at :361
    lib, st, dev = N.load(), N.stream_ptr(), t.device
    tb = _threshold_bits(H, W)
    out = torch.empty_like(t)
    bits = torch.empty((lib.svb_bits_words(H, W),), dtype=torch.int32, device=dev)
    rowcnt = torch.empty((H,), dtype=torch.int32, device=dev)
    status = torch.zeros((1,), dtype=torch.int32, device=dev)
    N.check(lib.svb_clean_road(N.ptr(t), N.ptr(tb), 1, H, W, N.ptr(out), N.ptr(bits), N.ptr(rowcnt), N.ptr(status), st),
            "svb_clean_road")
    if (height, width) != (H, W):
        # the reference scans only the top-left height x width window (:359-361)
        win = torch.zeros_like(out)
        win[:height, :width] = out[:height, :width]
        N.check(lib.svb_pack_mask(N.ptr(win), 1, H, W, N.ptr(bits), N.ptr(rowcnt), st), "svb_pack_mask")
        N.count(1)
    counts = torch.empty((1,), dtype=torch.int32, device=dev)
    N.check(lib.svb_mask_points(N.ptr(bits), N.ptr(rowcnt), 1, H, W, 0, 0, N.ptr(counts), st), "svb_mask_points")
    n = int(counts.item())
    if int(status.item()) & 1:
        raise ValueError("sanitiseRoadImage: the image holds values other than 0 and 255")
    pts = torch.empty((max(n, 1), 2), dtype=torch.int32, device=dev)
    N.check(lib.svb_mask_points(N.ptr(bits), N.ptr(rowcnt), 1, H, W, n, N.ptr(pts), N.ptr(counts), st), "svb_mask_points")
    N.count(3)
    # np.matrix([]) has shape (1,0) in the reference when nothing is left (:365)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", PendingDeprecationWarning)          # np.matrix is what the reference returns (:365)
        mat = np.matrix(pts[:n].cpu().numpy().astype(np.int64)) if n else np.matrix([])
    return _back(out, was_tensor), mat


# ---------------------------------------------------------------- f3 (obstacle stage)
def _hull_stage(t, masks, want_hull=True):
    lib, st, dev = N.load(), N.stream_ptr(), t.device
    H, W = t.shape
    bits = torch.empty((lib.svb_bits_words(H, W),), dtype=torch.int32, device=dev)
    N.check(lib.svb_pack_mask(N.ptr(t), 1, H, W, N.ptr(bits), None, st), "svb_pack_mask")
    cap = 2 * H
    hull = torch.empty((cap, 2), dtype=torch.int32, device=dev)
    meta = torch.empty((3,), dtype=torch.int32, device=dev)             # n, cx, cy
    hm = torch.empty_like(t) if masks else None
    ob = torch.empty_like(t) if masks else None
    sb = int(lib.svb_road_hull_scratch_bytes(1, H))
    scratch = torch.empty((sb,), dtype=torch.uint8, device=dev) if masks else None
    N.check(lib.svb_road_hull(N.ptr(bits), 1, H, W, cap, N.ptr(hull), N.ptr(meta), N.ptr(meta[1:]),
                              N.ptr(hm), N.ptr(ob), N.ptr(scratch), sb if masks else 0, st), "svb_road_hull")
    N.count(3 if masks else 2)
    n, cx, cy = (int(v) for v in meta.cpu())
    return hull[:n], (cx, cy), hm, ob


def roadObstacles(cleaned):
    """stereovision.py:170-180 as one call on the cleaned road image (0 / non-zero): returns
    (roadHull, hullMask, obstacleImage) -- cv2.convexHull of the road points as an (n,1,2) int32 array,
    drawContours(cleaned.copy(), [hull], 0, 255, -100), and NOT cleaned kept inside the hull mask.
    Raises ValueError for an empty image (the reference's convexHull fails on the empty matrix and the
    orchestrator skips the stage, stereovision.py:171,203)."""
    t, was_tensor = _to_dev_u8(cleaned)
    if t.dim() != 2:
        raise ValueError("roadObstacles expects a single-channel image")
    hull, _, hm, ob = _hull_stage(t, True)
    if hull.shape[0] == 0:
        raise ValueError("roadObstacles: no road points")
    h = hull.cpu().numpy().reshape(-1, 1, 2)
    return h, _back(hm, was_tensor), _back(ob, was_tensor)


def getCenterPoint(hull):
    """functions.py:418-421: int(x), int(y) of cv2.minEnclosingCircle(hull)'s centre.  The circle of a point
    set is the circle of its hull, so the points are drawn into a bit image and go through svb_road_hull."""
    p = np.asarray(hull.cpu() if torch.is_tensor(hull) else hull).reshape(-1, 2).astype(np.int64)
    if p.shape[0] == 0:
        raise ValueError("getCenterPoint: empty hull")
    if p.min() < 0 or p[:, 0].max() >= N.SVB_MAX_DIM or p[:, 1].max() >= N.SVB_MAX_DIM:
        raise ValueError("getCenterPoint: coordinates outside [0, %d)" % N.SVB_MAX_DIM)
    N.require_cuda()
    img = np.zeros((int(p[:, 1].max()) + 1, int(p[:, 0].max()) + 1), dtype=np.uint8)
    img[p[:, 1], p[:, 0]] = 255
    _, centre, _, _ = _hull_stage(torch.from_numpy(img).cuda(), False)
    return centre


# ---------------------------------------------------------------- f2
def _gamma_table(gamma):
    """functions.py:62-65, value by value in float64 (host side in the reference as well)."""
    inv = 1.0 / gamma
    return np.array([((i / 255.0) ** inv) * 255 for i in np.arange(0, 256)]).astype("uint8")


def gammaChange(image, gamma=1.0):
    """functions.py:61-67: every byte of the image through the gamma table (cv2.LUT)."""
    t, was_tensor = _to_dev_u8(image)
    table = np.ascontiguousarray(_gamma_table(gamma))
    lut = torch.from_numpy(table).to(t.device)
    out = torch.empty_like(t)
    if t.numel():
        N.check(N.load().svb_lut_bytes(N.ptr(t), t.numel(), N.ptr(lut), N.ptr(out), N.stream_ptr()), "svb_lut_bytes")
        N.count(1)
    return _back(out, was_tensor)


def preProcessImages(imgL, imgR):
    """functions.py:80-86: gamma 1.4 on both images."""
    return (gammaChange(imgL, 1.4), gammaChange(imgR, 1.4))


def greyscale(imgL, imgR):
    """functions.py:88-96: COLOR_BGR2GRAY then equalizeHist, per image."""
    lib, st, out = N.load(), N.stream_ptr(), []
    for img in (imgL, imgR):
        t, was_tensor = _to_dev_u8(img)
        if t.dim() != 3 or t.shape[2] != 3:
            raise ValueError("greyscale expects BGR images")
        H, W = int(t.shape[0]), int(t.shape[1])
        gray = torch.empty((H, W), dtype=torch.uint8, device=t.device)
        hist = torch.empty((256,), dtype=torch.int32, device=t.device)
        lut = torch.empty((256,), dtype=torch.uint8, device=t.device)
        N.check(lib.svb_gamma_gray_hist(N.ptr(t), 1, H, W, None, 0, N.ptr(gray), N.ptr(hist), st), "svb_gamma_gray_hist")
        N.check(lib.svb_equalize_lut(N.ptr(hist), 1, N.ptr(lut), st), "svb_equalize_lut")
        eq = torch.empty_like(gray)
        N.check(lib.svb_apply_lut_frames(N.ptr(gray), 1, H * W, N.ptr(lut), N.ptr(eq), st), "svb_apply_lut_frames")
        N.count(3)
        out.append(_back(eq, was_tensor))
    return (out[0], out[1])


# ---------------------------------------------------------------- f4
def disparity(grayL, grayR, max_disparity, crop_disparity):
    """functions.py:104-128 with the module's matcher StereoSGBM_create(0, 128, 21) (functions.py:26): semi-global
    matching + 3x3 median, filterSpeckles(.., 0, 4000, max_disparity - 5), threshold to zero, /16 -> uint8,
    optional crop [0:390, 135:], x (256 / max_disparity) -> uint8."""
    from .pipeline import Matcher
    L, was_tensor = _to_dev_u8(grayL)
    R, _ = _to_dev_u8(grayR)
    if L.dim() != 2 or L.shape != R.shape:
        raise ValueError("disparity expects two gray images of one shape")
    key = ("matcher", tuple(L.shape), int(max_disparity), str(L.device))
    if _state.get("matcher_key") != key:
        _state["matcher"] = Matcher(shape=tuple(L.shape), max_disparity=int(max_disparity), max_frames=1, frames_in_flight=1,
                                    device=L.device)
        _state["matcher_key"] = key
    out = _state["matcher"].disparity(L, R)[0]
    if int(_state["matcher"].status.item()) & 1:
        import warnings
        warnings.warn("disparity: a block cost exceeded 32767 - P2; cv2.StereoSGBM's own int16 stores wrap or saturate there "
                      "depending on its SIMD path, so this output is not specified", RuntimeWarning)
    if crop_disparity:
        out = out[0:390, 135:out.shape[1]]
    return _back(out.clone(), was_tensor)


# ---------------------------------------------------------------- f3
def fillAltDisparity(disparity):
    """functions.py:149-162: entries below 2 become the row's mean over its non-zero entries.  The reference
    writes into its argument and returns it; so does this (a numpy argument is updated in place)."""
    t, was_tensor = _to_dev_u8(disparity)
    if t.dim() != 2:
        raise ValueError("fillAltDisparity expects a single-channel image")
    H, W = t.shape
    out = t if was_tensor else torch.empty_like(t)
    N.check(N.load().svb_fill_row_mean(N.ptr(t), 1, H, W, N.ptr(out), N.stream_ptr()), "svb_fill_row_mean")
    N.count(1)
    if was_tensor:
        return out
    res = out.cpu().numpy()
    if isinstance(disparity, np.ndarray) and disparity.dtype == np.uint8 and disparity.flags.writeable:
        disparity[...] = res
        return disparity
    return res


def getNormalVectorLine(basePoint, abc, disparity):
    """functions.py:396-416: end point of the normal glyph for base point (x, y).  `disparity` is indexed
    [y, x] as in the reference; raises what the reference's int() conversions raise."""
    x, y = basePoint
    dv = int(disparity[y, x])                              # IndexError for a point outside, as in the reference
    dev = _dev()
    base = torch.tensor([[int(x), int(y)]], dtype=torch.int32, device=dev)
    pl = torch.from_numpy(np.ascontiguousarray(np.asarray(abc, dtype=np.float64).reshape(3))).to(dev)
    dval = torch.tensor([dv], dtype=torch.uint8, device=dev)
    out = torch.empty((1, 2), dtype=torch.int32, device=dev)
    status = torch.empty((1,), dtype=torch.int32, device=dev)
    cc = N.ccam(_camera())
    N.check(N.load().svb_normal_line(N.ptr(base), N.ptr(pl), 3, 0, N.ptr(dval), 1, 1, 1, C.byref(cc), N.ptr(out), N.ptr(status),
                                     N.stream_ptr()), "svb_normal_line")
    N.count(1)
    st = int(status.item())
    if st == 1:
        raise ValueError("cannot convert float NaN to integer")
    if st == 2:
        raise OverflowError("cannot convert float infinity to integer")
    o = out.cpu().numpy()
    return (int(o[0, 0]), int(o[0, 1]))


def getBlackImage():
    """functions.py:448-450."""
    return blackImg

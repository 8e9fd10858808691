"""TEST INFRASTRUCTURE -- CPU restatement (numpy, float64) of the reference's per-frame hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this file, and only as the checker or the timed CPU arm.  The product (stereo.vision_b200) never
does.  Parity pinning: every function below is checked against outputs of the *unmodified*
reference (functions.py executed in the build container by oracle/ref_harness.py) stored under
tests/golden/ -- see tests/test_oracle_golden.py.  The reference ships no tests of its own
(SURVEY 4), so those fixtures are the only pin there is.

Each function cites the reference lines it restates.  Arrays replace the reference's Python lists:
a cloud is (sites [N,2] int32 as (y,x), d [N] uint8, xyz [N,3] float64, rgb [N,3] uint8) in the
reference's append order (row-major over the lattice).
"""
import random

import numpy as np

F_PX = 399.9745178222656        # functions.py:15
BASE_M = 0.2090607502           # functions.py:19
CENTRE_H = 262.0                # functions.py:21
CENTRE_W = 474.5                # functions.py:22
HUE_BINS = 1001                 # str(round(h,3)) has 1001 possible values '0.0' .. '1.0'


class Cam:
    def __init__(self, f=F_PX, B=BASE_M, ch=CENTRE_H, cw=CENTRE_W):
        self.f, self.B, self.ch, self.cw = float(f), float(B), float(ch), float(cw)


def cam_of(camera=None):
    if camera is None:
        return Cam()
    if isinstance(camera, Cam):
        return camera
    return Cam(camera.focal_px, camera.baseline_m, camera.centre_h, camera.centre_w)


# ---------------------------------------------------------------- a1-a3: disparity clean / mask
def fill_disparity(d, prev):
    """functions.py:140-147: where d <= 2 add the previous frame's value, saturating at 255."""
    if prev is None:
        return d
    hole = d <= 2                                   # NOT(threshold(d, 2) -> 255 where d > 2)
    s = d.astype(np.uint16) + np.where(hole, prev, 0).astype(np.uint16)
    return np.minimum(s, 255).astype(np.uint8)      # cv2.add saturates


def mask_disparity(d, carmask):
    """functions.py:169-172: keep d where carmask != 0, else 0 (new array)."""
    return np.where(carmask != 0, d, 0).astype(np.uint8)


def cap_disparity(d):
    """functions.py:164-167: the bitwise_and result is discarded -- identity, same object."""
    return d


# ---------------------------------------------------------------- a4: projection
def lattice_shape(H, W, step=2):
    """Sites visited by range(0,H-1,step) x range(0,W-1,step) (functions.py:185-186)."""
    return len(range(0, H - 1, step)), len(range(0, W - 1, step))


def project(d, camera=None, step=2, bgr=None):
    """functions.py:178-198.  Returns dict(sites, d, xyz, rgb) in row-major lattice order.

    Z = (f*B)/d ; X = ((x-cw)*Z)/f ; Y = ((y-ch)*Z)/f, all float64, product/quotient order kept.
    rgb = bgr[y,x,2], bgr[y,x,1], bgr[y,x,0] (functions.py:195).
    """
    c = cam_of(camera)
    H, W = d.shape[:2]
    ys = np.arange(0, H - 1, step)
    xs = np.arange(0, W - 1, step)
    sub = d[np.ix_(ys, xs)]
    iy, ix = np.nonzero(sub > 0)                   # C order == the reference's append order
    y = ys[iy]
    x = xs[ix]
    dv = sub[iy, ix]
    fB = c.f * c.B
    Z = fB / dv.astype(np.float64)
    X = ((x.astype(np.float64) - c.cw) * Z) / c.f
    Y = ((y.astype(np.float64) - c.ch) * Z) / c.f
    out = {
        "sites": np.stack([y, x], axis=1).astype(np.int32),
        "d": dv.astype(np.uint8),
        "xyz": np.stack([X, Y, Z], axis=1),
        "rgb": None,
    }
    if bgr is not None and len(bgr) > 0:
        out["rgb"] = np.ascontiguousarray(bgr[y, x][:, ::-1]).astype(np.uint8)
    return out


# ---------------------------------------------------------------- a13: reprojection
def reproject(xyz, camera=None):
    """functions.py:201-209 then the caller's np.int32 truncation (stereovision.py:112).

    x = ((X*f)/Z) + cw ; y = ((Y*f)/Z) + ch in float64.  Returns (float [R,2], int32 [R,2]) as (x,y).
    """
    c = cam_of(camera)
    if len(xyz) == 0:
        z = np.zeros((0, 2))
        return z, z.astype(np.int32)
    X, Y, Z = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    x = ((X * c.f) / Z) + c.cw
    y = ((Y * c.f) / Z) + c.ch
    fl = np.stack([x, y], axis=1)
    return fl, fl.astype(np.int32)                  # astype truncates toward zero, as np.array(.., int32)


# ---------------------------------------------------------------- a10: hue key
def hue_bin(r, g, b):
    """functions.py:73-78 with colorsys.rgb_to_hsv fed np.uint8 channels.

    colorsys: maxc/minc; grey -> h = 0.0; rc,gc,bc = (maxc-ch)/(maxc-minc) (uint8/uint8 -> float64);
    h = bc-gc | 2+rc-bc | 4+gc-rc; h = (h/6) % 1 (np.remainder).  round(np.float64, 3) is
    rint(h*1000)/1000, so the dictionary key is identified by the integer rint(h*1000) in [0,1000].
    """
    r = np.asarray(r, dtype=np.uint8)
    g = np.asarray(g, dtype=np.uint8)
    b = np.asarray(b, dtype=np.uint8)
    maxc = np.maximum(np.maximum(r, g), b)
    minc = np.minimum(np.minimum(r, g), b)
    rng_ = (maxc - minc).astype(np.float64)
    grey = maxc == minc
    den = np.where(grey, 1.0, rng_)
    rc = (maxc - r).astype(np.float64) / den
    gc = (maxc - g).astype(np.float64) / den
    bc = (maxc - b).astype(np.float64) / den
    h = np.where(r == maxc, bc - gc, np.where(g == maxc, (2.0 + rc) - bc, (4.0 + gc) - rc))
    h = np.remainder(h / 6.0, 1.0)
    k = np.rint(h * 1000.0).astype(np.int32)
    return np.where(grey, 0, k).astype(np.int32)


def hue_key(bin_id: int) -> str:
    """The reference's dictionary key for a bin (functions.py:76-78)."""
    return str(round(np.float64(bin_id) / 1000.0, 3)) if bin_id else "0.0"


# ---------------------------------------------------------------- a5-a7: RANSAC
def replay_indices(mask_xyz, trials, n_samples, seed=None, rng=None):
    """The reference's draw order on index ranges (functions.py:286, 252-254).

    One random.sample(range(M), S) then rounds of three random.sample(range(M), 1) until the cross
    product of (P1-P2, P2-P3) is not exactly zero (functions.py:251-259).  M < S: every trial raises
    before drawing (functions.py:286,294) -> all indices -1.
    """
    if rng is None:
        rng = random.Random()
        rng.seed(seed)
    m = len(mask_xyz)
    samp = -np.ones((trials, n_samples), dtype=np.int32)
    trip = -np.ones((trials, 3), dtype=np.int32)
    if m < n_samples:
        return samp, trip
    pop = range(m)
    for t in range(trials):
        samp[t] = rng.sample(pop, n_samples)
        while True:
            i = [rng.sample(pop, 1)[0] for _ in range(3)]
            u = mask_xyz[i[0]] - mask_xyz[i[1]]
            v = mask_xyz[i[1]] - mask_xyz[i[2]]
            cp = (u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0])
            if not (cp[0] == 0 and cp[1] == 0 and cp[2] == 0):
                break
        trip[t] = i
    return samp, trip


def plane_of_triple(P):
    """functions.py:267: abc = inv([P1;P2;P3]) . ones.  Exactly singular -> None (trial skipped, :294)."""
    try:
        with np.errstate(all="ignore"):
            return np.dot(np.linalg.inv(P), np.ones(3))
    except np.linalg.LinAlgError:
        return None


def plane_distance(xyz, abc):
    """functions.py:269,275 / 307,310: |P.abc - 1| / sqrt(a^2+b^2+c^2), products summed left to right."""
    nrm = np.sqrt(abc[0] * abc[0] + abc[1] * abc[1] + abc[2] * abc[2])
    s = (xyz[:, 0] * abc[0] + xyz[:, 1] * abc[1]) + xyz[:, 2] * abc[2]
    return np.abs((s - 1.0) / nrm)


def image_collinear(sites3):
    """Three pixel sites on one image line: their plane contains the camera centre, [P1;P2;P3] has
    rank 2 in exact arithmetic and np.linalg.inv returns rounding noise (LinAlgError for about half
    of such triples here, |abc| ~ 1e14 otherwise -- machine dependent).  Both this oracle and the
    product treat the triple as singular (trial skipped, functions.py:294-296); measured on the
    golden frames such a hypothesis never comes near winning ("degenerate triples")."""
    (y1, x1), (y2, x2), (y3, x3) = [(int(a), int(b)) for a, b in sites3]
    return (x2 - x1) * (y3 - y1) - (x3 - x1) * (y2 - y1) == 0


def ransac_from_indices(mask_xyz, sample_idx, triple_idx, sites=None):
    """functions.py:278-298 with the draws given.  Returns (best_trial, abc or None, errors [T]).

    error = mean distance of the trial's S sample points to the trial's plane (:289); the first
    strict minimum wins (:291); a singular triple or a NaN error never wins.  `sites` [M,2] enables
    the image-collinear rule above.
    """
    T = len(sample_idx)
    errs = np.full(T, np.inf)
    planes = np.full((T, 3), np.nan)
    for t in range(T):
        if triple_idx[t, 0] < 0:
            continue
        if sites is not None and image_collinear(sites[triple_idx[t]]):
            continue
        abc = plane_of_triple(mask_xyz[triple_idx[t]])
        if abc is None:
            continue
        planes[t] = abc
        with np.errstate(all="ignore"):
            e = np.mean(plane_distance(mask_xyz[sample_idx[t]], abc))
        errs[t] = e if not np.isnan(e) else np.inf
    best, best_err = -1, np.inf
    for t in range(T):
        if errs[t] < best_err:
            best, best_err = t, errs[t]
    return best, (planes[best].copy() if best >= 0 else None), errs


# ---------------------------------------------------------------- a8-a14: road points
def road_points(cloud, abc, threshold=0.05, colour_threshold=10, shape=(544, 1024), camera=None):
    """stereovision.py:97-140: distances (functions.py:300-312), dist < threshold (:314-323), hue
    histogram (:215-226), count > colour_threshold (:228-230), reprojection (:201-209) truncated to
    int32 (stereovision.py:112), scatter of 255 into a black image (functions.py:339-344).
    """
    H, W = shape
    mask = np.zeros((H, W), dtype=np.uint8)
    out = {"near": None, "hist": np.zeros(HUE_BINS, dtype=np.int64), "keep": None,
           "near_n": 0, "road_n": 0, "pts": np.zeros((0, 2), dtype=np.int32), "mask": mask}
    if abc is None:
        return out
    dist = plane_distance(cloud["xyz"], abc)
    near = dist < threshold
    rgb = cloud["rgb"]
    bins = hue_bin(rgb[:, 0], rgb[:, 1], rgb[:, 2])
    hist = np.bincount(bins[near], minlength=HUE_BINS)
    keep = near & (hist[bins] > colour_threshold)
    _, pts = reproject(cloud["xyz"][keep], camera)
    mask[pts[:, 1], pts[:, 0]] = 255
    out.update(dist=dist, near=near, bins=bins, hist=hist, keep=keep, near_n=int(near.sum()),
               road_n=int(keep.sum()), pts=pts, mask=mask)
    return out


# ---------------------------------------------------------------- the frame, end to end
def run_frame(disparity, bgr, prev, carmask, seed=None, trials=600, n_samples=600, step=2,
              threshold=0.05, colour_threshold=10, camera=None, indices=None):
    """stereovision.py:58-140 on one frame.  `indices` = (sample_idx, triple_idx) overrides the replay."""
    filled = fill_disparity(disparity, prev)
    cloud = project(filled, camera, step, bgr)
    mcloud = project(mask_disparity(filled, carmask), camera, step)
    if indices is None:
        indices = replay_indices(mcloud["xyz"], trials, n_samples, seed)
    best, abc, errs = ransac_from_indices(mcloud["xyz"], indices[0], indices[1], mcloud["sites"])
    rp = road_points(cloud, abc, threshold, colour_threshold, disparity.shape[:2], camera)
    return {"filled": filled, "cloud": cloud, "mcloud": mcloud, "indices": indices, "best": best,
            "abc": abc, "errors": errs, "road": rp}


# ---------------------------------------------------------------- f1: road-mask cleaning
def _rect_morph(img, k, is_dilate):
    """cv2.dilate / cv2.erode with a k x k all-ones kernel, anchor at the centre and OpenCV's
    default border (morphologyDefaultBorderValue: pixels outside the image never win the max / min).
    Separable running max / min over the in-image part of the window."""
    r = k // 2
    H, W = img.shape
    op = np.maximum if is_dilate else np.minimum
    out = img.copy()
    for s in range(1, min(r, H - 1) + 1):           # rows: out[y] = op(img[y-r .. y+r])
        op(out[s:], img[:H - s], out=out[s:])
        op(out[:H - s], img[s:], out=out[:H - s])
    src = out
    out = src.copy()
    for s in range(1, min(r, W - 1) + 1):           # columns
        op(out[:, s:], src[:, :W - s], out=out[:, s:])
        op(out[:, :W - s], src[:, s:], out=out[:, :W - s])
    return out


def sanitise_road_image(img, road_threshold_mask):
    """functions.py:346-366 (sanitiseRoadImage) on a uint8 image; returns (cleaned image, pts).

    :350-351  closing with a 9x9 kernel (dilate then erode)
    :353      erode, 9x9, iterations=2
    :355      bitwise_and(img, img, mask=road_threshold_mask): kept where the mask is non-zero
    :356      closing again
    :358      ParticleCleansing (functions.py:378-384) paints every external contour whose area is
              below 70 px^2 *white*.  After a 9x9 closing that cannot change a pixel: the black set of
              a closed image is a union of 9x9 black squares, so a hole inside a white component holds
              at least 81 px and the contour around it has area >= 98 > 70; a component whose contour
              area is below 70 therefore has no hole, and painting the polygon through its own border
              pixels only touches pixels that are white already.  The golden fixtures, produced with
              the step left in, confirm it (tests/test_oracle_golden.py).
    :359-365  pts = [x, y] of every non-zero pixel, rows outermost (row-major), as an (N,2) matrix.
    """
    img = np.ascontiguousarray(img, dtype=np.uint8)
    a = _rect_morph(_rect_morph(img, 9, True), 9, False)
    a = _rect_morph(_rect_morph(a, 9, False), 9, False)
    a = np.where(np.asarray(road_threshold_mask) != 0, a, np.uint8(0))
    a = _rect_morph(_rect_morph(a, 9, True), 9, False)
    return a, nonzero_points(a)


def nonzero_points(img):
    """functions.py:359-365: (x, y) of the non-zero pixels in row-major order, int32 [N,2]."""
    ys, xs = np.nonzero(img)
    return np.stack([xs, ys], axis=1).astype(np.int32)


# ---------------------------------------------------------------- f2: pre-processing front end
def gamma_table(gamma=1.0):
    """functions.py:62-65: table[i] = uint8(((i / 255.0) ** (1 / gamma)) * 255), truncated, evaluated value by
    value in float64 as the reference's list comprehension does."""
    inv = 1.0 / gamma
    return np.array([((i / 255.0) ** inv) * 255 for i in np.arange(0, 256)]).astype("uint8")


def gamma_change(image, gamma=1.0):
    """functions.py:61-67 (cv2.LUT over every channel byte)."""
    return gamma_table(gamma)[np.asarray(image, dtype=np.uint8)]


def bgr_to_gray(bgr):
    """cv2.cvtColor(..., COLOR_BGR2GRAY) on uint8 (functions.py:93): OpenCV's fixed point,
    (B*3735 + G*19235 + R*9798 + 2^14) >> 15 (coefficients 0.114, 0.587, 0.299 in 15 bits)."""
    b = bgr[..., 0].astype(np.int32)
    g = bgr[..., 1].astype(np.int32)
    r = bgr[..., 2].astype(np.int32)
    return ((b * 3735 + g * 19235 + r * 9798 + (1 << 14)) >> 15).astype(np.uint8)


def equalize_lut(hist):
    """The look-up table of cv2.equalizeHist (functions.py:95) from a 256-bin histogram: first populated
    bin i0 maps to 0; lut[i] = saturate(rint(float32(sum_{i0<k<=i} hist[k]) * scale)) with the float32
    scale = 255 / (total - hist[i0]); an image of one value maps to itself."""
    hist = np.asarray(hist, dtype=np.int64)
    total = int(hist.sum())
    i0 = int(np.nonzero(hist)[0][0])
    if hist[i0] == total:
        return np.arange(256, dtype=np.uint8) * 0 + np.uint8(i0)
    scale = np.float32(255.0) / np.float32(total - hist[i0])
    s = np.cumsum(hist) - np.cumsum(hist)[i0]
    v = np.rint(s.astype(np.float32) * scale)
    lut = np.clip(v, 0, 255).astype(np.uint8)
    lut[:i0 + 1] = 0
    return lut


def equalize_hist(gray):
    gray = np.asarray(gray, dtype=np.uint8)
    return equalize_lut(np.bincount(gray.ravel(), minlength=256))[gray]


def preprocess_pair(imgL, imgR, gamma=1.4):
    """stereovision.py:44,46: gamma-corrected colour images (functions.py:80-86) and their equalised
    grey versions (functions.py:88-96)."""
    gl, gr = gamma_change(imgL, gamma), gamma_change(imgR, gamma)
    return gl, gr, equalize_hist(bgr_to_gray(gl)), equalize_hist(bgr_to_gray(gr))


# ---------------------------------------------------------------- f3: 'mean' fill and the normal glyph
def fill_alt_disparity(d):
    """functions.py:149-162 (threshold_option 'mean'): every row's entries below 2 become the row's mean over
    its non-zero entries (float64 mean, 0.0 for an empty row, truncated by the uint8 store).  Returns a new
    array; the reference writes into its argument."""
    d = np.array(d, dtype=np.uint8, copy=True)
    for i in range(d.shape[0]):
        row = d[i]
        nz = row[row != 0]
        mean = nz.mean() if len(nz) else 0.0
        row[row < 2] = np.uint8(mean)                 # C-style truncation, as the element-wise store does
    return d


def normal_vector_line(base_point, abc, d_at_base, camera=None):
    """functions.py:396-416 (getNormalVectorLine): the image point of the road point under `base_point`
    lifted 0.7 m along -Y and dropped back onto ... the reference's own formula, operation by operation in
    float64.  Returns (int, int) truncated toward zero; raises like int() does on nan / inf."""
    c = cam_of(camera)
    x, y = base_point
    a, b, cc = (np.float64(v) for v in np.asarray(abc, dtype=np.float64).reshape(3))
    with np.errstate(all="ignore"):
        Z = np.float64(c.f * c.B) / np.float64(d_at_base)
        X = ((x - c.cw) * Z) / c.f
        Y = ((y - c.ch) * Z) / c.f
        newY = Y - 0.7
        newX = X + 0.0
        dd = np.sqrt(a * a + b * b + cc * cc)
        Z = dd - ((a * newX) + (b * newY))
        newX = ((newX * c.f) / Z) + c.cw
        newY = ((newY * c.f) / Z) + c.ch
    return (int(newX), int(newY))


# ---------------------------------------------------------------- f3: convex hull, hull / obstacle masks, centre
def convex_hull(points):
    """cv2.convexHull of integer (x, y) points (stereovision.py:172, functions.py:392): the strict hull (no
    collinear vertices), in OpenCV's orientation, rotated to start at the lexicographically largest (x, y).
    OpenCV starts at a vertex of its own choosing (usually this one); callers compare up to rotation."""
    P = sorted(set((int(a), int(b)) for a, b in np.asarray(points).reshape(-1, 2)))
    if len(P) <= 1:
        return np.array(P, dtype=np.int32).reshape(-1, 2)

    def cross(o, a, b):
        return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])
    lo, up = [], []
    for p in P:
        while len(lo) >= 2 and cross(lo[-2], lo[-1], p) <= 0:
            lo.pop()
        lo.append(p)
    for p in reversed(P):
        while len(up) >= 2 and cross(up[-2], up[-1], p) <= 0:
            up.pop()
        up.append(p)
    return np.array(up[:-1] + lo[:-1], dtype=np.int32)


def _line_pixels(p1, p2):
    """cv::LineIterator, 8-connected, left to right (what drawContours uses for a polygon's edges)."""
    (x1, y1), (x2, y2) = p1, p2
    dx, dy = x2 - x1, y2 - y1
    if dx < 0:
        x1, y1, dx, dy = x2, y2, -dx, -dy
    ystep = 1
    if dy < 0:
        dy, ystep = -dy, -1
    out, x, y = [], x1, y1
    if dy > dx:
        err, plus, minus = dy - (dx + dx), dy + dy, -(dx + dx)
        for _ in range(dy + 1):
            out.append((x, y))
            if err < 0:
                err += minus + plus
                x += 1
            else:
                err += minus
            y += ystep
    else:
        err, plus, minus = dx - (dy + dy), dx + dx, -(dy + dy)
        for _ in range(dx + 1):
            out.append((x, y))
            if err < 0:
                err += minus + plus
                y += ystep
            else:
                err += minus
            x += 1
    return out


def hull_fill(hull, shape):
    """cv2.drawContours(zeros, [hull], 0, 255, -100) for a convex polygon (stereovision.py:176): OpenCV draws
    every edge with its 8-connected line iterator and scan-fills the interior, and for a convex polygon the
    union is, row by row, the span between the leftmost and the rightmost edge pixel (checked against cv2 on
    random hulls, segments and single points in the golden script)."""
    H, W = shape
    img = np.zeros((H, W), dtype=np.uint8)
    h = [tuple(int(v) for v in p) for p in np.asarray(hull).reshape(-1, 2)]
    if not h:
        return img
    lo, hi = np.full(H, 1 << 30), np.full(H, -1)
    for i in range(len(h)):
        for x, y in _line_pixels(h[i], h[(i + 1) % len(h)]):
            if 0 <= y < H:
                lo[y], hi[y] = min(lo[y], x), max(hi[y], x)
    for y in range(H):
        if hi[y] >= 0:
            img[y, max(lo[y], 0):min(hi[y], W - 1) + 1] = 255
    return img


def obstacle_stage(cleaned, points):
    """stereovision.py:172-180: hull of the road points, hull mask drawn over a copy of the cleaned image,
    obstacles = NOT cleaned, kept inside the hull mask.  Returns (hull, hull_mask, obstacle_mask)."""
    hull = convex_hull(points)
    fill = hull_fill(hull, cleaned.shape)
    hull_mask = np.where(fill != 0, np.uint8(255), cleaned)
    obstacles = np.where(hull_mask != 0, ~cleaned, np.uint8(0)).astype(np.uint8)
    return hull, hull_mask, obstacles


def enclosing_centre(hull):
    """functions.py:418-421 (getCenterPoint): int(x), int(y) of cv2.minEnclosingCircle's centre.  The circle
    itself is unique; OpenCV finds it in float32, this is the incremental algorithm in float64 on squared
    distances (equal after int() on every fixture; they can differ when a coordinate of the centre lies within
    ~1e-4 of an integer).  Every operation is a single rounded float64 one, so the CUDA kernel repeats it exactly."""
    P = [(float(x), float(y)) for x, y in np.asarray(hull).reshape(-1, 2)]

    def dist(ax, ay, bx, by):                    # squared
        dx, dy = ax - bx, ay - by
        return dx * dx + dy * dy

    def two(a, b):
        cx, cy = (a[0] + b[0]) / 2, (a[1] + b[1]) / 2
        return cx, cy, dist(a[0], a[1], cx, cy)

    def three(a, b, c):
        (ax, ay), (bx, by), (cx, cy) = a, b, c
        d = 2 * (ax * (by - cy) + bx * (cy - ay) + cx * (ay - by))
        if d == 0:
            return None
        ux = ((ax * ax + ay * ay) * (by - cy) + (bx * bx + by * by) * (cy - ay) + (cx * cx + cy * cy) * (ay - by)) / d
        uy = ((ax * ax + ay * ay) * (cx - bx) + (bx * bx + by * by) * (ax - cx) + (cx * cx + cy * cy) * (bx - ax)) / d
        return ux, uy, dist(ax, ay, ux, uy)

    def inside(c, p):
        return dist(p[0], p[1], c[0], c[1]) <= c[2] * (1 + 1e-11) + 1e-9      # c[2] = radius squared
    c = (P[0][0], P[0][1], 0.0)
    for i, p in enumerate(P):
        if inside(c, p):
            continue
        c = (p[0], p[1], 0.0)
        for j in range(i):
            if inside(c, P[j]):
                continue
            c = two(p, P[j])
            for k in range(j):
                if inside(c, P[k]):
                    continue
                cc = three(p, P[j], P[k])
                if cc:
                    c = cc
    return (int(c[0]), int(c[1]))

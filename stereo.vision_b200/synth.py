"""Seeded synthetic stereo frames for the path.

The reference's dataset is not in its repository (loop.py:3), so every test and benchmark feeds the
path the two arrays it actually consumes: the uint8 disparity that `f.disparity` would have produced
(functions.py:104-128, units of half pixels, 0 = no match) and the gamma-corrected BGR left image
(stereovision.py:44).  A frame is a pure function of (sequence id, frame id, shape); nothing here
touches the GPU and nothing is cached, so the GPU box regenerates exactly what the golden fixtures
were made from.
"""
import numpy as np

from .camera import DEFAULT_CAMERA, Camera


def frame_seed(seq: int, frame: int) -> int:
    return 1000 * int(seq) + int(frame)


def ransac_seed(seq: int, frame: int) -> int:
    """Seed handed to `random.seed` before the frame's RANSAC."""
    return 100003 * int(seq) + int(frame)


def true_plane(seq: int, frame: int):
    """(a, b, c) of the road plane a*X + b*Y + c*Z = 1 that make_frame renders for this frame"""
    rng = np.random.default_rng(frame_seed(seq, frame))
    jit = 1.0 + 0.05 * (rng.random(3) * 2.0 - 1.0)
    return np.array([0.02 * jit[0] * (1 if (seq + frame) % 2 == 0 else -1), (1.0 / 0.6) * jit[1], 0.03 * jit[2]])


def make_frame(seq: int, frame: int, shape=(544, 1024), camera: Camera = None, holes: float = 0.10):
    """Return (disparity uint8 [H,W], bgr uint8 [H,W,3]) of one synthetic frame.

    A tilted road plane a*X + b*Y + c*Z = 1 in the path's coordinates (Y down) is rendered to
    disparity d = f*B/Z; a few constant-disparity rectangles stand in for obstacles; +-1 integer
    noise and i.i.d. holes (d = 0) give the temporal fill something to do.  Road pixels are grey with
    a hue-coherent tint so that some hue bins collect more than `road_color_thresh` points and some
    do not; obstacles and sky are saturated random colours.
    """
    H, W = shape
    s = W / 1024.0
    cam = camera if camera is not None else (DEFAULT_CAMERA if s == 1.0 else DEFAULT_CAMERA.scaled(s))
    f, B, ch, cw = cam.focal_px, cam.baseline_m, cam.centre_h, cam.centre_w
    rng = np.random.default_rng(frame_seed(seq, frame))

    # road plane: camera ~0.6 m above the road, small roll/pitch, jittered per frame
    jit = 1.0 + 0.05 * (rng.random(3) * 2.0 - 1.0)
    a = 0.02 * jit[0] * (1 if (seq + frame) % 2 == 0 else -1)
    b = (1.0 / 0.6) * jit[1]
    c = 0.03 * jit[2]
    ys = (np.arange(H, dtype=np.float64) - ch)[:, None] / f
    xs = (np.arange(W, dtype=np.float64) - cw)[None, :] / f
    inv_z = a * xs + b * ys + c                       # 1/Z on the plane, <= 0 above the horizon
    d = np.rint(f * B * inv_z)
    road = d >= 1.0
    d = np.where(road, d, 0.0)

    d = d.astype(np.int16)

    def u8(*shape):
        return rng.integers(0, 256, size=shape, dtype=np.uint8)

    # sky / far field above the horizon: sparse weak matches (d in 1..5)
    far = (~road) & (u8(H, W) < 90)
    d = np.where(far, (u8(H, W) % 5 + 1).astype(np.int16), d)

    # obstacles: constant-disparity rectangles that drift slowly with the frame id
    n_obs = int(rng.integers(2, 7))
    obstacle = np.zeros((H, W), dtype=bool)
    orng = np.random.default_rng(7919 * int(seq) + 17)        # per-sequence layout
    for k in range(6):
        ow = int(orng.integers(40, 180) * s)
        oh = int(orng.integers(40, 160) * s)
        ox = int(orng.integers(0, W - ow))
        oy = int(orng.integers(int(H * 0.35), H - oh))
        od = int(orng.integers(20, 120))
        vx = int(orng.integers(-3, 4))
        if k >= n_obs:
            continue
        x0 = (ox + vx * frame) % (W - ow)
        d[oy:oy + oh, x0:x0 + ow] = od
        obstacle[oy:oy + oh, x0:x0 + ow] = True

    # matcher noise (+-1) and ~10 % holes
    noise = (u8(H, W) % 3).astype(np.int16) - 1
    d = np.where(d > 0, d + noise, 0)
    d = np.clip(d, 0, 254)
    d = np.where(u8(H, W) < int(round(holes * 256)), 0, d).astype(np.uint8)

    # colour: tinted grey road, saturated everything else
    grey = (u8(H, W) % 41).astype(np.int16) + 90
    tint = np.array([6, 2, -4], dtype=np.int16) + rng.integers(-1, 2, size=3).astype(np.int16)   # B, G, R
    wob = (u8(H, W, 3) % 7).astype(np.int16) - 3
    wob = np.where(u8(H, W, 1) < 205, wob // 2, wob * 3)     # a minority of off-hue road pixels
    road_bgr = np.clip(grey[:, :, None] + tint[None, None, :] + wob, 0, 255).astype(np.uint8)
    other_bgr = u8(H, W, 3)
    litter = u8(H, W) < 8                                      # ~3 % road pixels of arbitrary hue
    bgr = np.where((road & ~obstacle & ~litter)[:, :, None], road_bgr, other_bgr)
    return d, bgr


def make_sequence(seq: int, n_frames: int, shape=(544, 1024), camera: Camera = None, start: int = 0):
    """Return (disparity uint8 [F,H,W], bgr uint8 [F,H,W,3]) for frames start..start+n_frames-1."""
    H, W = shape
    disp = np.empty((n_frames, H, W), dtype=np.uint8)
    bgr = np.empty((n_frames, H, W, 3), dtype=np.uint8)
    for t in range(n_frames):
        disp[t], bgr[t] = make_frame(seq, start + t, shape, camera)
    return disp, bgr


def scale_mask(mask: np.ndarray, s: int) -> np.ndarray:
    """Nearest-neighbour upscale of a shipped 544x1024 mask (stress config)."""
    if s == 1:
        return mask
    return np.repeat(np.repeat(mask, s, axis=0), s, axis=1)


def _box3(a: np.ndarray) -> np.ndarray:
    """One 3x3 box blur with edge replication (float32)."""
    p = np.pad(a, 1, mode="edge")
    h = p[:, :-2] + p[:, 1:-1] + p[:, 2:]
    return (h[:-2] + h[1:-1] + h[2:]) / np.float32(9.0)


def make_stereo_pair(seq: int, frame: int, shape=(544, 1024), camera: Camera = None):
    """Return (grayL, grayR) uint8 [H,W]: a rectified pair whose true disparity is `make_frame`'s scene.

    Input of the matcher stage (functions.py:104-128).  The right view is a random texture
    (blurred noise plus a coarse pattern, so that both the gradient and the raw term of the matching cost
    carry signal); the left view looks it up at x - disparity (disparity in pixels = half of make_frame's
    units); pixels without a match (make_frame's holes) are independent noise in the left view, which is
    what gives the speckle filter and the left-right check work to do.
    """
    H, W = shape
    d, _ = make_frame(seq, frame, shape, camera)
    rng = np.random.default_rng(frame_seed(seq, frame) + 500009)
    fine = rng.integers(0, 256, size=(H, W), dtype=np.uint8).astype(np.float32)
    coarse = rng.integers(0, 256, size=((H + 7) // 8, (W + 7) // 8), dtype=np.uint8).astype(np.float32)
    coarse = np.repeat(np.repeat(coarse, 8, axis=0), 8, axis=1)[:H, :W]
    tex = 0.55 * _box3(_box3(fine)) + 0.45 * _box3(_box3(_box3(coarse)))
    tex = (tex - tex.min()) / max(float(tex.max() - tex.min()), 1.0) * 255.0
    right = tex.astype(np.uint8)
    px = (d.astype(np.int32) + 1) // 2
    xs = np.arange(W, dtype=np.int32)[None, :] - px
    left = np.take_along_axis(right, np.clip(xs, 0, W - 1), axis=1)
    lost = (d == 0) | (xs < 0)
    left = np.where(lost, rng.integers(0, 256, size=(H, W), dtype=np.uint8), left)
    glare = rng.integers(-2, 3, size=(H, W)).astype(np.int16)          # sensor noise on one view
    left = np.clip(left.astype(np.int16) + glare, 0, 255).astype(np.uint8)
    return np.ascontiguousarray(left), np.ascontiguousarray(right)


def make_stereo_bgr(seq: int, frame: int, shape=(544, 1024), camera: Camera = None):
    """Return (imgL, imgR) uint8 [H,W,3] BGR: `make_stereo_pair`'s views with a fixed tint (B+6, G+2, R-4), i.e. what
    the frame orchestrator takes (stereovision.py:26) -- camera images before gamma correction and grey conversion.
    A constant tint keeps the road's hue coherent, so most near-plane points survive the colour filter."""
    gl, gr = make_stereo_pair(seq, frame, shape, camera)
    tint = np.array([6, 2, -4], dtype=np.int16)[None, None, :]
    out = []
    for g in (gl, gr):
        out.append(np.clip(g.astype(np.int16)[:, :, None] + tint, 0, 255).astype(np.uint8))
    return np.ascontiguousarray(out[0]), np.ascontiguousarray(out[1])

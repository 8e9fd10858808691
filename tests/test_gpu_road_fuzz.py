"""Fuzz of the float32 decision band of road_sites_near (csrc/svb_sites.cu): the sites path decides `near` with one
float32 FMA chain per lattice site and hands the sites within E of the boundary to the reference's float64 sequence.
E is an analytic bound; here it is held to the exact records path (svb_road_points) and to the oracle over 576 planes
that a RANSAC run would rarely choose: random directions with |abc| from 1e-3 to 1e6, steep planes, perturbed road
planes, and planes built so that the threshold boundary |P.abc - 1| / |abc| = 0.05 passes through a lattice row's
points to within a few ulps (functions.py:300-323)."""
import ctypes as C

import numpy as np
import pytest

import np_oracle as no

pytestmark = pytest.mark.gpu

H, W, STEP, THR, CTHR = 96, 160, 2, 0.05, 3


BASE = np.array([0.02, -0.9, 0.25])      # the plane the test image is rendered from (rows 0-95 lie above the principal point)


def _image(rng, cam):
    """disparity of the plane BASE with +-1 noise, holes and a flat patch; colours with a few populous hues"""
    yy, xx = np.mgrid[0:H, 0:W]
    inv_z = BASE[0] * (xx - cam.cw) / cam.f + BASE[1] * (yy - cam.ch) / cam.f + BASE[2]
    d = np.clip(np.rint(cam.f * cam.B * inv_z) + rng.integers(-1, 2, (H, W)), 0, 254).astype(np.uint8)
    d[rng.random((H, W)) < 0.1] = 0
    d[20:40, 30:70] = 57
    bgr = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    bgr[:, : W // 2] = (bgr[:, : W // 2] // 64) * 64 + 20
    return d, bgr


def _boundary_planes(rng, cam, n, d_img):
    """planes a = 0 whose distance to the points of lattice row y0 at disparity d0 is THR up to a few ulps"""
    out = []
    fB = cam.f * cam.B
    while len(out) < n:
        y0 = int(rng.integers(0, (H - 1) // STEP)) * STEP
        row = d_img[y0, 0:W - 1:STEP]
        row = row[row > 0]
        d0 = int(np.bincount(row).argmax())                 # the disparity most sites of that row have
        c = float(rng.uniform(-0.2, 0.2))
        sign = 1.0 if rng.random() < 0.5 else -1.0
        Z0 = fB / d0
        k = (y0 - cam.ch) / cam.f

        def g(b):
            return abs(Z0 * (k * b + c) - 1.0) - THR * np.sqrt(b * b + c * c)
        lo, hi = (0.0, 50.0)
        grid = np.linspace(lo, hi, 4001) * sign
        vals = np.array([g(b) for b in grid])
        idx = np.nonzero(np.sign(vals[:-1]) * np.sign(vals[1:]) < 0)[0]
        if len(idx) == 0:
            continue
        i = int(rng.choice(idx))
        a_, b_ = grid[i], grid[i + 1]
        for _ in range(80):
            m = 0.5 * (a_ + b_)
            if np.sign(g(m)) == np.sign(g(a_)):
                a_ = m
            else:
                b_ = m
        b = 0.5 * (a_ + b_)
        for ulps in (-2, 0, 3):
            out.append(np.array([0.0, b * (1.0 + ulps * 2.0 ** -52), c]))
    return out[:n]


def _planes(rng, cam, d_img):
    pl = []
    for _ in range(200):                                   # any direction, |abc| over nine decades
        v = rng.normal(size=3)
        pl.append(v / np.linalg.norm(v) * 10.0 ** rng.uniform(-3, 6))
    for _ in range(120):                                   # road-like planes, perturbed
        pl.append(BASE * (1.0 + 0.03 * rng.normal(size=3)))
    for _ in range(64):                                    # steep: dominated by a or by c
        v = np.array([rng.normal() * 5, rng.normal() * 0.05, rng.normal() * 0.05]) if rng.random() < 0.5 else \
            np.array([rng.normal() * 0.01, rng.normal() * 0.01, 10.0 ** rng.uniform(-2, 1)])
        pl.append(v)
    pl += _boundary_planes(rng, cam, 192, d_img)
    return np.array(pl)


def test_sites_path_decision_band_against_records_path_and_oracle(cuda):
    import torch
    import stereo.vision_b200._native as N
    from stereo.vision_b200.pipeline import RoadPath
    lib = N.load()
    rng = np.random.default_rng(2024)
    cam = no.cam_of(None)
    d, bgr = _image(rng, cam)
    planes = _planes(rng, cam, d)
    assert len(planes) == 576
    cm = np.full((H, W), 255, np.uint8)
    cloud_o = no.project(d, None, STEP, bgr)
    FB = 64
    eng = RoadPath(shape=(H, W), step=STEP, trials=8, samples=8, carmask=cm, max_frames=FB, point_threshold=THR, colour_threshold=CTHR)
    rec = RoadPath(shape=(H, W), step=STEP, trials=8, samples=8, carmask=cm, max_frames=FB, point_threshold=THR, colour_threshold=CTHR,
                   road_path="records")
    assert eng.sites_road and not rec.sites_road
    dd = np.broadcast_to(d, (FB, H, W)).copy()
    cc = np.broadcast_to(bgr, (FB, H, W, 3)).copy()
    eng.process(dd, cc, seed=1, fill=False)                  # leaves filled / cloud / counts of FB identical frames
    rec.process(dd, cc, seed=1, fill=False)
    bgr_d = torch.from_numpy(cc).to(cuda)
    near_total = unsure_like = 0
    for base in range(0, len(planes), FB):
        rows = np.zeros((FB, 8))
        for i in range(FB):
            abc = np.ascontiguousarray(planes[base + i])
            row = np.zeros(8)
            assert lib.svb_plane_from_abc(abc.ctypes.data_as(C.c_void_p), THR, row.ctypes.data_as(C.c_void_p)) == 0
            assert row[7] == 1.0
            rows[i] = row
        pd = torch.from_numpy(rows).to(cuda)
        out = {}
        for name, e in (("sites", eng), ("records", rec)):
            hist = torch.zeros((FB, 1024), dtype=torch.int32, device=cuda)
            mask = torch.full((FB, H, W), 7, dtype=torch.uint8, device=cuda)
            counts = e.counts[:FB].clone()
            if name == "sites":
                N.check(lib.svb_road_points_sites(N.ptr(e.filled), N.ptr(bgr_d), N.ptr(pd), FB, H, W, STEP, C.byref(e._cam), CTHR,
                                                  N.ptr(e.reproj_bits), N.ptr(hist), N.ptr(mask), None, N.ptr(counts),
                                                  N.ptr(e.scratch), e.scratch.numel(), N.stream_ptr()), "sites")
            else:
                N.check(lib.svb_road_points(N.ptr(e.cloud), N.ptr(e.counts), N.ptr(pd), FB, H, W, STEP, C.byref(e._cam), CTHR,
                                            N.ptr(e.reproj_bits), N.ptr(hist), N.ptr(mask), None, N.ptr(counts),
                                            N.ptr(e.scratch), e.scratch.numel(), N.stream_ptr()), "records")
            out[name] = (mask.cpu().numpy(), hist.cpu().numpy(), counts.cpu().numpy())
        for k in range(3):                                   # masks, hue bins (the sites path keeps a counter behind them), counts
            a, b = out["sites"][k], out["records"][k]
            assert np.array_equal(a[:, :no.HUE_BINS], b[:, :no.HUE_BINS]) if k == 1 else np.array_equal(a, b), ("sites vs records", base, k)
        for i in range(FB):
            o = no.road_points(cloud_o, planes[base + i], THR, CTHR, (H, W))
            m, h, c = (out["sites"][k][i] for k in range(3))
            assert (int(c[2]), int(c[3])) == (o["near_n"], o["road_n"]), (base + i, planes[base + i], c, o["near_n"], o["road_n"])
            assert np.array_equal(h[:no.HUE_BINS].astype(np.int64), o["hist"]) and np.array_equal(m, o["mask"]), base + i
            near_total += o["near_n"]
            # how close the plane's boundary comes to a point: the fuzz is only worth something if it comes close often
            dist = o["dist"] if "dist" in o else None
            if dist is not None and len(dist):
                unsure_like += int((np.abs(dist - THR) < 1e-9).sum())
    assert near_total > 500000 and unsure_like > 3000, (near_total, unsure_like)

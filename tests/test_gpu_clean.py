"""GPU: road-mask cleaning (sanitiseRoadImage functions.py:346-366) through the C-ABI
(svb_pack_mask / svb_clean_road / svb_clean_road_bits / svb_mask_points) against the unmodified
reference's outputs (tests/golden/sanitise.npz) and the numpy oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

import np_oracle as no
import stereo.vision_b200._native as N
from helpers import RedZone, sanitise_cases
from stereo.vision_b200.masks import default_road_threshold_mask
from stereo.vision_b200.pipeline import RoadCleaner
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


def bit_rows(img):
    """numpy restatement of the bit-row layout of include/svb200.h: uint32 [H, ceil(W/32)], LSB first."""
    H, W = img.shape
    WW = (W + 31) // 32
    pad = np.zeros((H, WW * 32), np.uint8)
    pad[:, :W] = img != 0
    return np.packbits(pad, axis=1, bitorder="little").view("<u4").reshape(H, WW)


def run_clean(cuda, imgs, tm, cap=None, packed=False, want_bytes=True):
    lib = N.load()
    imgs = np.ascontiguousarray(imgs)
    F, H, W = imgs.shape
    words = lib.svb_bits_words(H, W)
    d_in = torch.from_numpy(imgs).to(cuda)
    d_tm = torch.from_numpy(np.ascontiguousarray(tm)).to(cuda)
    tb = torch.empty((words,), dtype=torch.int32, device=cuda)
    N.check(lib.svb_pack_mask(N.ptr(d_tm), 1, H, W, N.ptr(tb), 0, 0), "pack")
    rz = RedZone(cuda)
    out = rz.alloc((F, H, W), torch.uint8, 7) if want_bytes else None
    bits = rz.alloc((F, words), torch.int32)
    rowcnt = rz.alloc((F, H), torch.int32)
    status = rz.alloc((1,), torch.int32, -1)
    if packed:
        ib = torch.empty((F, words), dtype=torch.int32, device=cuda)
        N.check(lib.svb_pack_mask(N.ptr(d_in), F, H, W, N.ptr(ib), 0, 0), "pack")
        N.check(lib.svb_clean_road_bits(N.ptr(ib), N.ptr(tb), F, H, W, N.ptr(out) if want_bytes else 0, N.ptr(bits), N.ptr(rowcnt), 0),
                "svb_clean_road_bits")
        status.zero_()
    else:
        N.check(lib.svb_clean_road(N.ptr(d_in), N.ptr(tb), F, H, W, N.ptr(out) if want_bytes else 0, N.ptr(bits), N.ptr(rowcnt),
                                   N.ptr(status), 0), "svb_clean_road")
    counts = torch.empty((F,), dtype=torch.int32, device=cuda)
    N.check(lib.svb_mask_points(N.ptr(bits), N.ptr(rowcnt), F, H, W, 0, 0, N.ptr(counts), 0), "svb_mask_points")
    cap = int(counts.max().item()) if cap is None else cap
    pts = rz.alloc((F, max(cap, 1), 2), torch.int32, -5)
    N.check(lib.svb_mask_points(N.ptr(bits), N.ptr(rowcnt), F, H, W, cap, N.ptr(pts), N.ptr(counts), 0), "svb_mask_points")
    torch.cuda.synchronize()
    rz.check()
    return {"out": out.cpu().numpy() if want_bytes else None, "bits": bits.cpu().numpy().view(np.uint32).reshape(F, H, -1),
            "rowcnt": rowcnt.cpu().numpy(), "counts": counts.cpu().numpy(), "pts": pts.cpu().numpy(), "status": int(status.item())}


def check_against_oracle(res, imgs, tm):
    for f in range(len(imgs)):
        want, wpts = no.sanitise_road_image(imgs[f], tm)
        if res["out"] is not None:
            assert np.array_equal(res["out"][f], want), "frame %d cleaned image" % f
        assert np.array_equal(res["bits"][f], bit_rows(want)), "frame %d bit rows" % f
        assert np.array_equal(res["rowcnt"][f], (want != 0).sum(axis=1))
        n = int(res["counts"][f])
        assert n == len(wpts)
        assert np.array_equal(res["pts"][f, :n], wpts), "frame %d point order" % f


@pytest.fixture(params=[4, 2, 1, 0], ids=lambda b: "bands%d" % b)
def bands(request):
    """row bands per frame of the 1024-wide kernel; 0 = the general kernel"""
    N.check(N.load().svb_debug_clean_bands(request.param), "bands")
    yield request.param
    N.check(N.load().svb_debug_clean_bands(4), "bands")


def test_golden_cases(cuda, bands):
    """every case of the reference fixture: image sha, point count, point list sha -- bit exact"""
    cases = sanitise_cases()
    full = [(n, i, t, r) for n, i, t, r in cases if i.shape == (544, 1024)]
    imgs = np.stack([i for _, i, _, _ in full])
    res = run_clean(cuda, imgs, full[0][2])
    assert res["status"] == 0
    for k, (name, _, _, r) in enumerate(full):
        assert sha(res["out"][k]) == str(r["out_sha"]), name
        n = int(res["counts"][k])
        assert n == int(r["n_pts"]), name
        assert sha(res["pts"][k, :n]) == str(r["pts_sha"]), name
    name, img, tm, r = [c for c in cases if c[1].shape != (544, 1024)][0]
    res = run_clean(cuda, img[None], tm)                      # 70 x 101: ragged last word, scalar loads
    assert sha(res["out"][0]) == str(r["out_sha"])
    assert np.array_equal(res["pts"][0, :int(res["counts"][0])], r["pts"])


@pytest.mark.parametrize("shape", [(2, 2), (9, 33), (25, 31), (64, 96), (130, 257), (544, 1024), (300, 2048), (1088, 2048), (2048, 48)])
def test_shapes_against_oracle(cuda, shape):
    """band splits, halos clipped by the image, ragged rows, the largest frame (config 5's 2048 x 1088)"""
    H, W = shape
    rng = np.random.default_rng(H * 4099 + W)
    F = 3 if H * W < 1 << 21 else 1
    imgs = ((rng.random((F, H, W)) < 0.3) * 255).astype(np.uint8)
    imgs[:, H // 3:, W // 8: W - W // 8] |= ((rng.random((F, H - H // 3, W - W // 8 - W // 8)) < 0.5) * 255).astype(np.uint8)
    imgs[0, :, :] |= ((rng.random((H, W)) < 0.9) * 255).astype(np.uint8)          # nearly full: erosion from the borders only
    tm = ((rng.random((H, W)) < 0.97) * 128).astype(np.uint8)
    tm[: H // 5] = 0
    res = run_clean(cuda, imgs, tm)
    assert res["status"] == 0
    check_against_oracle(res, imgs, tm)
    res2 = run_clean(cuda, imgs, tm, packed=True, want_bytes=False)
    assert np.array_equal(res2["bits"], res["bits"]) and np.array_equal(res2["pts"], res["pts"])


def test_empty_full_and_flagged_inputs(cuda, bands):
    tm = default_road_threshold_mask()
    imgs = np.zeros((3, 544, 1024), np.uint8)
    imgs[1] = 255
    imgs[2, 300:400, 300:600] = 200                    # not a 0/255 image: flagged, treated as set
    res = run_clean(cuda, imgs, tm)
    assert res["status"] == 1
    assert res["counts"][0] == 0 and not res["out"][0].any()
    assert np.array_equal(res["out"][1], np.where(tm != 0, 255, 0))         # erosion never bites from the border
    b = imgs[2].copy()
    b[b != 0] = 255
    assert np.array_equal(res["out"][2], no.sanitise_road_image(b, tm)[0])
    assert run_clean(cuda, imgs[:2], tm)["status"] == 0


def test_point_capacity_and_scatter_roundtrip(cuda):
    """counts report the true size when cap is short; scatter(points(mask)) == mask (a14 is the inverse)"""
    name, img, tm, r = sanitise_cases()[1]
    res = run_clean(cuda, img[None], tm, cap=1000)
    n = int(r["n_pts"])
    assert int(res["counts"][0]) == n and res["pts"].shape[1] == 1000
    full = run_clean(cuda, img[None], tm)
    assert np.array_equal(res["pts"][0], full["pts"][0, :1000])
    lib = N.load()
    p = torch.from_numpy(full["pts"][0, :n].copy()).to(cuda)
    canvas = torch.zeros((544, 1024), dtype=torch.uint8, device=cuda)
    oob = torch.zeros((1,), dtype=torch.int32, device=cuda)
    N.check(lib.svb_scatter_points(N.ptr(p), n, N.ptr(canvas), 544, 1024, 1, (C.c_uint8 * 4)(255, 0, 0, 0), N.ptr(oob), 0), "scatter")
    assert int(oob.item()) == 0 and np.array_equal(canvas.cpu().numpy(), full["out"][0])


def test_pack_mask_layout(cuda):
    rng = np.random.default_rng(3)
    lib = N.load()
    for H, W in ((5, 7), (17, 64), (33, 100), (544, 1024)):
        img = rng.integers(0, 3, (2, H, W), dtype=np.uint8) * np.uint8(100)
        d = torch.from_numpy(img).to(cuda)
        bits = torch.empty((2, lib.svb_bits_words(H, W)), dtype=torch.int32, device=cuda)
        rc = torch.empty((2, H), dtype=torch.int32, device=cuda)
        N.check(lib.svb_pack_mask(N.ptr(d), 2, H, W, N.ptr(bits), N.ptr(rc), 0), "pack")
        got = bits.cpu().numpy().view(np.uint32).reshape(2, H, -1)
        for f in range(2):
            assert np.array_equal(got[f], bit_rows(img[f]))
            assert np.array_equal(rc[f].cpu().numpy(), (img[f] != 0).sum(axis=1))


def test_argument_errors(cuda):
    lib = N.load()
    x = torch.zeros((544, 1024), dtype=torch.uint8, device=cuda)
    tb = torch.zeros((lib.svb_bits_words(544, 1024),), dtype=torch.int32, device=cuda)
    assert lib.svb_clean_road(0, N.ptr(tb), 1, 544, 1024, N.ptr(x), 0, 0, 0, 0) == N.SVB_E_NULL
    assert lib.svb_clean_road(N.ptr(x), N.ptr(tb), 1, 544, 1024, 0, 0, 0, 0, 0) == N.SVB_E_NULL       # no output at all
    assert lib.svb_clean_road(N.ptr(x), N.ptr(tb), 1, 544, 1024, N.ptr(x), 0, 0, 0, 0) == N.SVB_E_ARG     # in place
    assert lib.svb_clean_road(N.ptr(x), N.ptr(tb), 1, 4096, 1024, N.ptr(x), 0, 0, 0, 0) == N.SVB_E_SHAPE
    assert lib.svb_mask_points(N.ptr(tb), 0, 1, 544, 1024, 0, 0, 0, 0) == N.SVB_E_NULL


def test_batch_of_300_frames_is_frame_independent(cuda):
    """BASELINE config 2 size: every frame of a 300-frame call equals the same frame cleaned alone, and
    the cleaned set obeys the operator's invariants (inside the threshold mask, closed under closing)."""
    rng = np.random.default_rng(11)
    base = [c[1] for c in sanitise_cases() if c[1].shape == (544, 1024)]
    tm = default_road_threshold_mask()
    cl = RoadCleaner(max_frames=300, device=cuda)
    masks = torch.empty((300, 544, 1024), dtype=torch.uint8, device=cuda)
    pick = rng.integers(0, len(base), 300)
    shift = rng.integers(-40, 40, 300)
    host = {}
    for f in range(300):
        img = np.roll(base[pick[f]], int(shift[f]), axis=1)
        masks[f] = torch.from_numpy(img).to(cuda)
        if f % 37 == 0:
            host[f] = img
    cleaned, counts, points = cl.clean(masks)
    torch.cuda.synchronize()
    assert int(cl.status.item()) == 0
    tmd = torch.from_numpy(tm).to(cuda)
    assert not bool((cleaned[:, tmd == 0] != 0).any())
    assert torch.equal((cleaned != 0).sum(dim=(1, 2)).to(torch.int32), counts)
    for f, img in host.items():
        want, wpts = no.sanitise_road_image(img, tm)
        assert np.array_equal(cleaned[f].cpu().numpy(), want)
        assert np.array_equal(points[f, :int(counts[f])].cpu().numpy(), wpts)


def test_dropin_sanitise_road_image(cuda):
    """the helper as stereovision.py:156 calls it: (image, np.matrix of [x, y] rows)"""
    import stereo.vision_b200.functions as f
    name, img, tm, r = sanitise_cases()[0]
    out, pts = f.sanitiseRoadImage(img, (544, 1024))
    assert isinstance(pts, np.matrix) and pts.shape == (int(r["n_pts"]), 2)
    assert sha(out) == str(r["out_sha"]) and sha(np.asarray(pts, dtype=np.int32)) == str(r["pts_sha"])
    out, pts = f.sanitiseRoadImage(f.getBlackImage(), (544, 1024))          # stereovision.py:160 feeds black frames too
    assert not out.any() and pts.shape == (1, 0)
    out, pts = f.sanitiseRoadImage(img, (400, 1024))                         # the scan window is `size`
    want, wpts = no.sanitise_road_image(img, tm)
    assert np.array_equal(out, want) and np.array_equal(np.asarray(pts), wpts[wpts[:, 1] < 400])
    grey = img.copy()
    grey[0, 0] = 9
    with pytest.raises(ValueError):
        f.sanitiseRoadImage(grey, (544, 1024))
    name, simg, stm, sr = sanitise_cases()[-1]                               # own threshold mask (module global, :33)
    old = f.road_threshold_mask
    try:
        f.road_threshold_mask = stm
        out, pts = f.sanitiseRoadImage(simg, simg.shape)
        assert sha(out) == str(sr["out_sha"]) and np.array_equal(np.asarray(pts), sr["pts"])
    finally:
        f.road_threshold_mask = old

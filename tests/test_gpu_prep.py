"""GPU: the front end (gammaChange, BGR->gray, equalizeHist; functions.py:61-67, 80-96) through
the C-ABI against the unmodified reference's outputs (tests/golden/preprocess.npz) and the numpy oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

import np_oracle as no
import stereo.vision_b200._native as N
from helpers import RedZone, preprocess_cases
from stereo.vision_b200.pipeline import FrontEnd
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


def run_front(cuda, imgs, table):
    """imgs uint8 [F,H,W,3] -> gamma bgr, gray, hist, lut, equalised gray (numpy) via the three entry points"""
    lib = N.load()
    imgs = np.ascontiguousarray(imgs)
    F, H, W, _ = imgs.shape
    d = torch.from_numpy(imgs).to(cuda)
    rz = RedZone(cuda)
    out = rz.alloc((F, H, W, 3), torch.uint8, 9)
    gray = rz.alloc((F, H, W), torch.uint8, 9)
    hist = rz.alloc((F, 256), torch.int32, -1)
    lut = rz.alloc((F, 256), torch.uint8)
    eq = rz.alloc((F, H, W), torch.uint8)
    tp = None if table is None else np.ascontiguousarray(table).ctypes.data_as(C.c_void_p)
    N.check(lib.svb_gamma_gray_hist(N.ptr(d), F, H, W, tp, N.ptr(out), N.ptr(gray), N.ptr(hist), 0), "svb_gamma_gray_hist")
    N.check(lib.svb_equalize_lut(N.ptr(hist), F, N.ptr(lut), 0), "svb_equalize_lut")
    N.check(lib.svb_apply_lut_frames(N.ptr(gray), F, H * W, N.ptr(lut), N.ptr(eq), 0), "svb_apply_lut_frames")
    torch.cuda.synchronize()
    rz.check()
    return [t.cpu().numpy() for t in (out, gray, hist, lut, eq)]


def test_golden_pairs(cuda):
    """left / right images of every fixture case: gamma-corrected colours and equalised gray, bit exact"""
    cases, g = preprocess_cases()
    table = no.gamma_table(1.4)
    assert np.array_equal(table, g["gamma_tables"][0])
    for name, imgL, imgR, rec in cases:
        out, gray, hist, lut, eq = run_front(cuda, np.stack([imgL, imgR]), table)
        assert sha(out[0]) == str(rec["gammaL_sha"]) and sha(out[1]) == str(rec["gammaR_sha"]), name
        assert sha(eq[0]) == str(rec["grayL_sha"]) and sha(eq[1]) == str(rec["grayR_sha"]), name
        assert np.array_equal(np.bincount(eq[0].ravel(), minlength=256), rec["grayL_hist"]), name
        assert np.array_equal(hist[0], np.bincount(gray[0].ravel(), minlength=256)), name


def test_gray_of_all_colours(cuda):
    """BGR2GRAY over all 2^24 colours against cv2's result stored in the fixture"""
    _, g = preprocess_cases()
    r, gg, b = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
    cube = np.stack([b, gg, r], axis=-1).reshape(1, 4096, 4096, 3)
    lib = N.load()
    d = torch.from_numpy(np.ascontiguousarray(cube)).to(cuda)
    gray = torch.empty((4, 2048, 2048), dtype=torch.uint8, device=cuda)              # the cube as four 2048 x 2048 images
    N.check(lib.svb_gamma_gray_hist(N.ptr(d), 4, 2048, 2048, None, 0, N.ptr(gray), 0, 0), "svb_gamma_gray_hist")
    got = gray.cpu().numpy().reshape(4096, 4096)
    assert sha(got) == str(g["gray_all_sha"])


@pytest.mark.parametrize("shape", [(1, 1), (3, 5), (37, 53), (64, 96), (544, 1024), (1088, 2048)])
def test_shapes_against_oracle(cuda, shape):
    H, W = shape
    rng = np.random.default_rng(H * 31 + W)
    F = 3
    imgs = rng.integers(0, 256, (F, H, W, 3), dtype=np.uint8)
    imgs[1] = (imgs[1] // 32) * 3 + 90                    # few levels
    imgs[2] = 200                                         # one level: equalizeHist maps the image to itself
    table = no.gamma_table(2.2)
    out, gray, hist, lut, eq = run_front(cuda, imgs, table)
    for f in range(F):
        gl = no.gamma_change(imgs[f], 2.2)
        y = no.bgr_to_gray(gl)
        assert np.array_equal(out[f], gl) and np.array_equal(gray[f], y)
        h = np.bincount(y.ravel(), minlength=256)
        assert np.array_equal(hist[f], h)
        assert np.array_equal(lut[f][h > 0], no.equalize_lut(h)[h > 0])
        assert np.array_equal(eq[f], no.equalize_hist(y))
    out2, gray2, _, _, _ = run_front(cuda, imgs, None)      # no table: identity colours
    assert np.array_equal(out2, imgs) and np.array_equal(gray2[0], no.bgr_to_gray(imgs[0]))


def test_front_end_batch_and_dropin(cuda):
    """FrontEnd on a batch equals the per-image drop-in helpers, which equal the reference's outputs"""
    import stereo.vision_b200.functions as f
    cases, g = preprocess_cases()
    name, imgL, imgR, rec = cases[0]
    gl, gr = f.preProcessImages(imgL, imgR)
    yl, yr = f.greyscale(gl, gr)
    assert sha(gl) == str(rec["gammaL_sha"]) and sha(gr) == str(rec["gammaR_sha"])
    assert sha(yl) == str(rec["grayL_sha"]) and sha(yr) == str(rec["grayR_sha"])
    for k, gamma in enumerate(g["gammas"]):
        ramp = np.arange(256, dtype=np.uint8).reshape(16, 16)
        assert np.array_equal(f.gammaChange(ramp, float(gamma)).reshape(256), g["gamma_tables"][k])
    fe = FrontEnd(max_frames=6, device=cuda)
    batch = np.stack([cases[0][1], cases[0][2], cases[1][1], cases[1][2], cases[2][1], cases[3][1]])
    bgr, eq = fe.process(batch)
    torch.cuda.synchronize()
    assert sha(bgr[0].cpu().numpy()) == str(rec["gammaL_sha"]) and sha(eq[1].cpu().numpy()) == str(rec["grayR_sha"])
    assert sha(eq[2].cpu().numpy()) == str(cases[1][3]["grayL_sha"]) and sha(eq[4].cpu().numpy()) == str(cases[2][3]["grayL_sha"])
    assert sha(eq[5].cpu().numpy()) == str(cases[3][3]["grayL_sha"])


def test_argument_errors(cuda):
    lib = N.load()
    x = torch.zeros((8, 8, 3), dtype=torch.uint8, device=cuda)
    assert lib.svb_gamma_gray_hist(0, 1, 8, 8, None, N.ptr(x), 0, 0, 0) == N.SVB_E_NULL
    assert lib.svb_gamma_gray_hist(N.ptr(x), 1, 8, 8, None, 0, 0, 0, 0) == N.SVB_E_NULL
    assert lib.svb_gamma_gray_hist(N.ptr(x), 0, 8, 8, None, N.ptr(x), 0, 0, 0) == N.SVB_E_SHAPE
    assert lib.svb_equalize_lut(0, 1, N.ptr(x), 0) == N.SVB_E_NULL

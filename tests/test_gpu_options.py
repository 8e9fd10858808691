"""GPU: SURVEY 8(f)-3 -- fillAltDisparity ('mean' hole filling, functions.py:149-162) and getNormalVectorLine
(functions.py:396-416) through the C-ABI against the unmodified reference (tests/golden/options.npz) and the oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

import np_oracle as no
import stereo.vision_b200._native as N
from helpers import RedZone, option_cases
from stereo.vision_b200.camera import DEFAULT_CAMERA
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


def test_fill_row_mean_golden_and_shapes(cuda):
    lib = N.load()
    imgs, g = option_cases()
    for i, im in enumerate(imgs):
        d = torch.from_numpy(im).to(cuda)
        out = torch.full_like(d, 7)
        N.check(lib.svb_fill_row_mean(N.ptr(d), 1, im.shape[0], im.shape[1], N.ptr(out), 0), "svb_fill_row_mean")
        assert sha(out.cpu().numpy()) == str(g["fill%d_out_sha" % i]), i
        N.check(lib.svb_fill_row_mean(N.ptr(d), 1, im.shape[0], im.shape[1], N.ptr(d), 0), "svb_fill_row_mean")     # in place
        assert torch.equal(d, out)
    rng = np.random.default_rng(1)
    for H, W in ((1, 1), (5, 17), (33, 100), (64, 2048), (300, 1024)):
        batch = (rng.integers(0, 256, (3, H, W), dtype=np.uint8) * (rng.random((3, H, W)) < 0.5)).astype(np.uint8)
        batch[1, :, :] = rng.integers(0, 3, (H, W), dtype=np.uint8)
        d = torch.from_numpy(batch).to(cuda)
        rz = RedZone(cuda)
        out = rz.alloc((3, H, W), torch.uint8)
        N.check(lib.svb_fill_row_mean(N.ptr(d), 3, H, W, N.ptr(out), 0), "svb_fill_row_mean")
        torch.cuda.synchronize()
        rz.check()
        got = out.cpu().numpy()
        for f in range(3):
            assert np.array_equal(got[f], no.fill_alt_disparity(batch[f])), (H, W, f)


def test_normal_line_golden(cuda):
    lib = N.load()
    _, g = option_cases()
    planes, cases, results = g["normal_planes"], g["normal_cases"], g["normal_results"]
    n = len(cases)
    base = torch.from_numpy(np.ascontiguousarray(cases[:, 1:3].astype(np.int32))).to(cuda)
    abc = torch.from_numpy(np.ascontiguousarray(planes[cases[:, 0]])).to(cuda)
    dv = torch.from_numpy(np.ascontiguousarray(cases[:, 3].astype(np.uint8))).to(cuda)
    out = torch.empty((n, 2), dtype=torch.int32, device=cuda)
    st = torch.empty((n,), dtype=torch.int32, device=cuda)
    cc = N.ccam(DEFAULT_CAMERA)
    N.check(lib.svb_normal_line(N.ptr(base), N.ptr(abc), 3, 0, N.ptr(dv), n, 1, 1, C.byref(cc), N.ptr(out), N.ptr(st), 0), "svb_normal_line")
    out, st = out.cpu().numpy(), st.cpu().numpy()
    for i in range(n):
        want = results[i]
        if want[0] == -(2 ** 31):
            assert st[i] == want[1], cases[i]
        else:
            assert st[i] == 0 and tuple(out[i]) == tuple(want), cases[i]
    # the same through an image: value read at [y, x]; a point outside reports status 4
    img = torch.zeros((1, 544, 1024), dtype=torch.uint8, device=cuda)
    img[0, 300, 512] = 40
    b1 = torch.tensor([[512, 300]], dtype=torch.int32, device=cuda)
    o1 = torch.empty((1, 2), dtype=torch.int32, device=cuda)
    s1 = torch.empty((1,), dtype=torch.int32, device=cuda)
    N.check(lib.svb_normal_line(N.ptr(b1), N.ptr(abc), 3, N.ptr(img), 0, 1, 544, 1024, C.byref(cc), N.ptr(o1), N.ptr(s1), 0), "svb_normal_line")
    assert tuple(o1.cpu().numpy()[0]) == no.normal_vector_line((512, 300), planes[0], 40) and int(s1.item()) == 0
    b1[0, 0] = 2000
    N.check(lib.svb_normal_line(N.ptr(b1), N.ptr(abc), 3, N.ptr(img), 0, 1, 544, 1024, C.byref(cc), N.ptr(o1), N.ptr(s1), 0), "svb_normal_line")
    assert int(s1.item()) == 4


def test_dropins(cuda):
    import stereo.vision_b200.functions as f
    imgs, g = option_cases()
    d = imgs[2].copy()
    r = f.disparityCleaning(d, 'mean', None)                 # stereovision.py:58 with threshold_option 'mean'
    assert r is d and sha(d) == str(g["fill2_out_sha"])      # the reference writes into its argument
    assert sha(f.fillAltDisparity(imgs[0].copy())) == str(g["fill0_out_sha"])
    planes = g["normal_planes"]
    disp = np.zeros((544, 1024), np.uint8)
    disp[300, 512] = 40
    assert f.getNormalVectorLine((512, 300), planes[0].reshape(3, 1), disp) == no.normal_vector_line((512, 300), planes[0], 40)
    with pytest.raises((ValueError, OverflowError)):
        f.getNormalVectorLine((0, 0), planes[0].reshape(3, 1), disp)      # disparity 0 under the base point
    with pytest.raises(IndexError):
        f.getNormalVectorLine((5000, 0), planes[0].reshape(3, 1), disp)

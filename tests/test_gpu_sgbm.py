"""GPU: the matcher stage (functions.disparity = cv2.StereoSGBM_create(0,128,21).compute + 3x3 median,
cv2.filterSpeckles, scaling; functions.py:26,104-128) through the C-ABI against the unmodified reference's outputs
(tests/golden/sgbm.npz) and the C oracle (oracle/sgbm_oracle.c, pinned to the same fixture on the CPU)."""
import numpy as np
import pytest
import torch

import sgbm_oracle as so
import stereo.vision_b200._native as N
import stereo.vision_b200.functions as sf
import stereo.vision_b200.synth as synth
from helpers import RedZone, sgbm_golden, sgbm_inputs
from stereo.vision_b200.pipeline import Matcher
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


def gpu_compute(cuda, L, R, block=21, frames_in_flight=None):
    """svb_sgbm_compute on [F,H,W] pairs with sentinel margins around output and scratch"""
    lib = N.load()
    L, R = np.ascontiguousarray(L), np.ascontiguousarray(R)
    if L.ndim == 2:
        L, R = L[None], R[None]
    F, H, W = L.shape
    rz = RedZone(cuda)
    out = rz.alloc((F, H, W), torch.int16, 77)
    nb = lib.svb_sgbm_scratch_bytes(frames_in_flight or F, H, W, 128)
    scratch = rz.alloc((nb,), torch.uint8)
    dL, dR = torch.from_numpy(L).to(cuda), torch.from_numpy(R).to(cuda)          # named: both must stay alive
    N.check(lib.svb_sgbm_compute(N.ptr(dL), N.ptr(dR), F, H, W, 128, block, N.ptr(out), None, N.ptr(scratch), nb, 0), "svb_sgbm_compute")
    torch.cuda.synchronize()
    rz.check()
    return out.cpu().numpy()


def gpu_speckles(cuda, img16, new_val, max_size, max_diff):
    lib = N.load()
    img16 = np.ascontiguousarray(img16)
    if img16.ndim == 2:
        img16 = img16[None]
    F, H, W = img16.shape
    rz = RedZone(cuda)
    d = rz.alloc((F, H, W), torch.int16)
    d.copy_(torch.from_numpy(img16))
    nb = lib.svb_speckle_scratch_bytes(F, H, W)
    scratch = rz.alloc((nb,), torch.uint8)
    N.check(lib.svb_filter_speckles(N.ptr(d), F, H, W, new_val, max_size, max_diff, N.ptr(scratch), nb, 0), "svb_filter_speckles")
    torch.cuda.synchronize()
    rz.check()
    return d.cpu().numpy()


@pytest.mark.parametrize("case", [c[0] for c in sgbm_inputs()])
def test_golden_case(cuda, case):
    """raw matcher output, both speckle settings and the final uint8 disparity of every fixture case, bit exact"""
    g = sgbm_golden()
    name, L, R, crop = next(c for c in sgbm_inputs() if c[0] == case)
    p = name + "_"
    raw = gpu_compute(cuda, L, R)[0]
    assert int((raw >= 0).sum()) == int(g[p + "valid"])
    assert sha(raw) == str(g[p + "raw_sha"])
    assert sha(gpu_speckles(cuda, raw, 0, 60, 40)[0]) == str(g[p + "speckle60_sha"])
    spk = gpu_speckles(cuda, raw, 0, 4000, 123)[0]
    assert int(((raw != 0) & (spk == 0)).sum()) == int(g[p + "despeckled"])
    assert sha(spk) == str(g[p + "speckle_sha"])
    final = sf.disparity(L, R, 128, crop)
    assert final.dtype == np.uint8 and final.shape == tuple(g[p + "shape"])
    assert int((final != 0).sum()) == int(g[p + "nonzero"])
    assert sha(final) == str(g[p + "final_sha"])
    if p + "final" in g:
        assert np.array_equal(final, g[p + "final"])


@pytest.mark.parametrize("block", [1, 5, 9, 31])
def test_block_sizes_against_oracle(cuda, block):
    L, R = synth.make_stereo_pair(2, 5, (120, 400))
    L, R = L[:90, :301], R[:90, :301]
    want = so.sgbm_compute(L, R, 128, block)
    got = gpu_compute(cuda, L, R, block=block)[0]
    assert np.array_equal(got, want)


def test_stress_frame_2048x1088(cuda):
    """BASELINE config 5's frame size: one 2048x1088 pair (more shared memory per row than the default limit,
    1.07 GB of cost volume) against the oracle"""
    L, R = synth.make_stereo_pair(5, 0, (1088, 2048))
    got = gpu_compute(cuda, L, R)[0]
    assert np.array_equal(got, so.sgbm_compute(L, R))
    assert np.array_equal(gpu_speckles(cuda, got, 0, 4000, 123)[0], so.filter_speckles(got.copy(), 0, 4000, 123))


@pytest.mark.parametrize("shift", [0, 5, 37, 100])
def test_shifted_pair_known_answer(cuda, shift):
    """size-independent property at the full frame size: the right view is the left view moved by `shift` pixels, so
    the block cost at d = shift is exactly zero wherever the window stays inside both views; the winner is `shift`
    there and the sub-pixel term (at most half a pixel) is all that can move the x16 value; shift 0 gives exactly 0."""
    rng = np.random.default_rng(31 + shift)
    H, W = 544, 1024
    tex = rng.integers(0, 256, (H, W + 128), dtype=np.uint8).astype(np.float32)
    tex = (tex + np.roll(tex, 1, 1) + np.roll(tex, 1, 0) + np.roll(np.roll(tex, 1, 0), 1, 1)) / 4
    tex = tex.astype(np.uint8)
    L = np.ascontiguousarray(tex[:, :W])
    R = np.ascontiguousarray(tex[:, shift:shift + W])                 # R[x] = L[x + shift]: left pixel x matches right pixel x - shift
    raw = gpu_compute(cuda, np.stack([L, L]), np.stack([R, R]))
    assert np.array_equal(raw[0], raw[1]), "frames of a batch are independent"
    inner = raw[0][16:-16, 128 + 16:W - 16].astype(np.int32)
    assert (inner >= 0).all()
    if shift == 0:
        assert (inner == 0).all()
    else:
        assert np.abs(inner - 16 * shift).max() <= 8
    d8 = sf.disparity(L, R, 128, False)
    assert set(np.unique(d8[16:-16, 128 + 16:W - 16]).tolist()) <= {2 * shift - 2, 2 * shift}


def test_batch_in_chunks(cuda):
    """five pairs with room for two frames of cost volume: three chunks, every frame equal to its single run"""
    pairs = [synth.make_stereo_pair(9, t, (120, 400)) for t in range(5)]
    L = np.stack([p[0] for p in pairs])
    R = np.stack([p[1] for p in pairs])
    got = gpu_compute(cuda, L, R, frames_in_flight=2)
    for t in range(5):
        assert np.array_equal(got[t], so.sgbm_compute(L[t], R[t])), "frame %d" % t
    m = Matcher(shape=(120, 400), max_frames=5, frames_in_flight=3, device=cuda)
    d8 = m.disparity(L, R).cpu().numpy()
    for t in range(5):
        assert np.array_equal(d8[t], so.disparity(L[t], R[t], 128, False)), "frame %d" % t


def test_speckles_random_labels(cuda):
    """filterSpeckles alone: random piecewise-constant images, every size / difference setting against the oracle"""
    rng = np.random.default_rng(5)
    for H, W, levels, max_size, max_diff in [(64, 96, 4, 20, 0), (100, 333, 3, 150, 16), (37, 41, 6, 5, 1), (200, 512, 2, 4000, 123),
                                             (1, 300, 3, 4, 0), (300, 1, 3, 4, 0), (128, 128, 2, 100000, 0)]:
        coarse = rng.integers(0, levels, ((H + 3) // 4, (W + 5) // 6))
        img = np.repeat(np.repeat(coarse, 4, axis=0), 6, axis=1)[:H, :W]
        img = (img * 16 - 16 + rng.integers(0, 2, (H, W))).astype(np.int16)          # -16, 0 (= newVal) and small steps
        want = so.filter_speckles(img.copy(), 0, max_size, max_diff)
        got = gpu_speckles(cuda, img, 0, max_size, max_diff)[0]
        assert np.array_equal(got, want), (H, W, levels, max_size, max_diff)
    # long snakes: one component winding through the whole image
    img = np.zeros((129, 200), np.int16)
    img[::2, :] = 48
    img[1::4, -1] = 48
    img[3::4, 0] = 48
    for max_size in (100, 20000):
        assert np.array_equal(gpu_speckles(cuda, img, 0, max_size, 0)[0], so.filter_speckles(img.copy(), 0, max_size, 0))


def test_scale_all_values(cuda):
    """functions.py:117-127 over every value the matcher can emit (and all negatives), for the reference's
    max_disparity = 128 and for 256 (no uint8 wrap in either)"""
    lib = N.load()
    v = np.arange(-32768, 2048, dtype=np.int16)
    d = torch.from_numpy(v).to(cuda)
    for maxd in (128, 256):
        out = torch.empty((v.size,), dtype=torch.uint8, device=cuda)
        N.check(lib.svb_disparity_scale(N.ptr(d), v.size, maxd, N.ptr(out), 0), "svb_disparity_scale")
        want = so.scale_disparity(v[None].copy(), maxd, False)[0]
        assert np.array_equal(out.cpu().numpy(), want), maxd


def test_argument_errors(cuda):
    lib = N.load()
    t = torch.zeros((1, 40, 200), dtype=torch.uint8, device=cuda)
    o = torch.zeros((1, 40, 200), dtype=torch.int16, device=cuda)
    s = torch.zeros((lib.svb_sgbm_scratch_bytes(1, 40, 200, 128),), dtype=torch.uint8, device=cuda)
    assert lib.svb_sgbm_compute(None, N.ptr(t), 1, 40, 200, 128, 21, N.ptr(o), None, N.ptr(s), s.numel(), 0) == N.SVB_E_NULL
    assert lib.svb_sgbm_compute(N.ptr(t), N.ptr(t), 1, 40, 200, 64, 21, N.ptr(o), None, N.ptr(s), s.numel(), 0) == N.SVB_E_ARG
    assert lib.svb_sgbm_compute(N.ptr(t), N.ptr(t), 1, 40, 200, 128, 20, N.ptr(o), None, N.ptr(s), s.numel(), 0) == N.SVB_E_ARG
    assert lib.svb_sgbm_compute(N.ptr(t), N.ptr(t), 1, 40, 130, 128, 21, N.ptr(o), None, N.ptr(s), s.numel(), 0) == N.SVB_E_ARG
    assert lib.svb_sgbm_compute(N.ptr(t), N.ptr(t), 1, 40, 200, 128, 21, N.ptr(o), None, N.ptr(s), 1000, 0) == N.SVB_E_SCRATCH
    assert lib.svb_filter_speckles(N.ptr(o), 1, 40, 200, 0, 10, 1, N.ptr(s), 100, 0) == N.SVB_E_SCRATCH
    with pytest.raises(ValueError):
        sf.disparity(np.zeros((40, 200), np.uint8), np.zeros((41, 200), np.uint8), 128, False)


def test_block_cost_overflow_raises_the_status_bit(cuda):
    """a block cost above 32767 - P2 (where cv2's own int16 stores wrap or saturate depending on its SIMD path) is
    reported in status bit 0; ordinary pairs leave it clear.  At the reference's block size (21 x 21 = 441 pixels) the bound
    is a mean pixel cost of 74, which 8-bit images hardly reach (the half-pixel interval of the Birchfield-Tomasi cost
    absorbs edges; the raw term wants flat areas, the gradient term slopes); at the largest supported block (31 x 31) it is
    34, and wide stripes against their inverse exceed it (so do ordinary pairs at wrong disparities)."""
    import warnings
    import stereo.vision_b200.functions as f
    from stereo.vision_b200.pipeline import Matcher
    stripes = np.tile(np.repeat(np.array([0, 255], np.uint8), 8), 19)[:300]
    hi = np.ascontiguousarray(np.broadcast_to(stripes, (64, 300)))
    mt = Matcher(shape=(64, 300), block=31, max_frames=1, frames_in_flight=1, device=cuda)
    mt.disparity(hi, (255 - hi).astype(np.uint8))
    assert int(mt.status.item()) & 1
    mt21 = Matcher(shape=(64, 300), block=21, max_frames=1, frames_in_flight=1, device=cuda)
    mt21.disparity(hi, (255 - hi).astype(np.uint8))
    assert int(mt21.status.item()) == 0                                   # 441 pixels: the same input stays below the bound
    gl, gr = synth.make_stereo_pair(3, 1, (120, 400))
    mt2 = Matcher(shape=(120, 400), block=21, max_frames=1, frames_in_flight=1, device=cuda)      # the reference's block size
    mt2.disparity(gl, gr)
    assert int(mt2.status.item()) == 0
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        f.disparity(gl, gr, 128, False)
    assert not any(issubclass(x.category, RuntimeWarning) for x in w)

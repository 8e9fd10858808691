"""GPU: the drop-in helpers (stereo.vision_b200.functions) called in the orchestrator's order
(stereovision.py:58-140) against outputs of the unmodified reference (tests/golden)."""
import random

import numpy as np
import pytest

import np_oracle as no
import stereo.vision_b200.synth as synth
from test_oracle_golden import frame, load, sha

pytestmark = pytest.mark.gpu


def run_chain(f, d, bgr, prev, seed, trials, thr, cthr):
    """The calls of stereovision.py:58-140, written the way the orchestrator writes them."""
    out = {}
    disparity = f.disparityCleaning(d, 'previous', prev)
    out["filled"] = np.asarray(disparity)
    maskedDisparity = f.maskDisparity(disparity)
    cappedDisparity = f.capDisparity(disparity)
    assert cappedDisparity is disparity
    points = f.projectDisparityTo3d(cappedDisparity, 128, bgr)
    maskpoints = f.projectDisparityTo3d(maskedDisparity, 128)
    out["points"], out["maskpoints"] = points, maskpoints
    random.seed(seed)
    normal, abc = f.RANSAC(maskpoints, trials)
    out["abc"] = abc
    out["rng_after"] = random.random()
    if abc is None:
        with pytest.raises(TypeError):
            f.calculatePointErrors(abc, points)
        out["mask"] = f.generatePointsAsImage(np.zeros((0, 1, 2), np.int32))
        return out
    assert normal is abc and abc.shape == (3, 1) and abc.dtype == np.float64
    diffs = f.calculatePointErrors(abc, points)
    points = f.computePlanarThreshold(points, diffs, thr)
    out["near_n"] = len(points)
    histogram = f.calculateColourHistogram(points)
    points = f.filterPointsByHistogram(points, histogram, cthr)
    out["road_n"] = len(points)
    planePoints = f.project3DPointsTo2DImagePoints(points)
    planePoints = np.array(planePoints, np.int32)
    planePoints = planePoints.reshape((-1, 1, 2))
    out.update(diffs=np.asarray(diffs), hist=histogram, pts=planePoints, mask=f.generatePointsAsImage(planePoints))
    return out


def check(out, r):
    assert sha(out["filled"]) == str(r["filled_sha"])
    assert len(out["points"]) == int(r["N"]) and len(out["maskpoints"]) == int(r["M"])
    assert sha(out["points"].xyz) == str(r["xyz_sha"])               # float64 XYZ bit-identical
    assert sha(out["points"].rgb) == str(r["rgb_sha"])
    assert sha(out["maskpoints"].xyz) == str(r["mask_xyz_sha"])
    if not bool(r["has_plane"]):
        assert out["abc"] is None and not out["mask"].any()
        return
    np.testing.assert_allclose(out["abc"].reshape(3), r["abc"], rtol=1e-9)
    assert out["near_n"] == int(r["near_n"]) and out["road_n"] == int(r["road_n"])
    bins = np.zeros(no.HUE_BINS, np.int64)
    for k, v in out["hist"].items():
        bins[int(round(float(k) * 1000))] = v
    assert np.array_equal(bins, r["hist_bins"])
    assert sha(out["pts"].reshape(-1, 2)) == str(r["pts_sha"])
    assert sha(out["mask"]) == str(r["road_mask_sha"])
    np.testing.assert_allclose(out["diffs"].reshape(-1)[:256], r["diffs_head"], rtol=1e-10, atol=1e-13)


@pytest.fixture()
def f(cuda):
    import importlib
    import stereo.vision_b200.functions as mod
    mod = importlib.reload(mod)
    from stereo.vision_b200.masks import default_carmask
    mod.carmask = default_carmask((544, 1024), prefer_cwd=False)
    yield mod
    importlib.reload(mod)                     # the tests rebind module settings (masks, step, sample size): leave none behind


def test_reference_sequence_s600(f):
    g = load("seq0_step2_s600.npz")
    prev = None
    for t in range(int(g["n_frames"])):
        d, bgr = synth.make_frame(0, t)
        seed = synth.ransac_seed(0, t)
        out = run_chain(f, d, bgr, prev, seed, 600, 0.05, 10)
        prev = out["filled"]
        check(out, frame(g, t))
        # the global `random` stream is left exactly where the reference leaves it
        rng = random.Random()
        rng.seed(seed)
        no.replay_indices(no.project(no.mask_disparity(out["filled"], f.carmask))["xyz"], 600, 600, rng=rng)
        assert out["rng_after"] == rng.random()


def test_reference_s400_and_step1(f):
    g = load("seq6_step2_s400.npz")
    f.ransac_samples = 400
    d, bgr = synth.make_frame(6, 0)
    check(run_chain(f, d, bgr, None, synth.ransac_seed(6, 0), 600, 0.05, 10), frame(g, 0))
    g = load("seq7_step1_s600.npz")
    f.ransac_samples, f.lattice_step = 600, 1
    d, bgr = synth.make_frame(7, 0)
    check(run_chain(f, d, bgr, None, synth.ransac_seed(7, 0), 600, 0.05, 10), frame(g, 0))


def test_small_frame_and_no_plane(f):
    g = load("small_64x96.npz")
    r = frame(g, 0)
    f.carmask, f.blackImg = r["in_carmask"], np.zeros((64, 96), np.uint8)
    out = run_chain(f, r["in_disp"], r["in_bgr"], r["in_prev"], int(g["seed"]), int(g["trials"]),
                    float(g["point_threshold"]), int(g["colour_threshold"]))
    check(out, r)
    assert np.array_equal(np.asarray(out["points"])[:, :3], r["xyz"])
    rows = list(out["points"])
    assert len(rows) == len(r["xyz"]) and rows[3][:3] == list(r["xyz"][3]) and tuple(rows[3][3:]) == tuple(r["rgb"][3])
    g = load("noplane.npz")
    f.carmask = np.zeros((544, 1024), np.uint8)
    f.carmask[200:210, 100:160] = 255
    f.blackImg = np.zeros((544, 1024), np.uint8)
    d, bgr = synth.make_frame(2, 0)
    check(run_chain(f, d, bgr, None, 1, 600, 0.05, 10), frame(g, 0))

"""GPU: whole frames.  tests/frame_harness.performStereoVision (the reference's stage sequence on the drop-in
helpers, every stage on the GPU) against tests/golden/frames.npz, which oracle/make_golden_frame.py recorded from the
UNMODIFIED reference orchestrator (stereovision.py:26-247) on three chained synthetic BGR stereo frames: matcher
disparity, carried disparity, plane, point counts, road image, cleaned image and points, hull centre, normal line."""
import os
import random

import numpy as np
import pytest

import frame_harness as sv
import stereo.vision_b200.synth as synth
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def fresh_helpers():
    """the helper module carries rebindable settings (masks, lattice step, sample size): start from its defaults"""
    import importlib
    import stereo.vision_b200.functions as mod
    importlib.reload(mod)
    yield


GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "frames.npz")


def test_three_chained_frames(cuda):
    with np.load(GOLD, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    seq = int(g["seq"])
    opts = dict(sv.default_opts)
    opts['loop'] = False
    prev = None
    for t in range(int(g["n_frames"])):
        p = "f%d_" % t
        imgL, imgR = synth.make_stereo_bgr(seq, t)
        assert sha(imgL) + sha(imgR) == str(g[p + "in_sha"]), "synthetic input drifted"
        random.seed(synth.ransac_seed(seq, t))
        images, prev, normal = sv.performStereoVision(imgL.copy(), imgR.copy(), prev, opts)
        st = images["stats"]
        assert st["Frame"] == t + 1
        assert sha(np.asarray(prev)) == str(g[p + "filled_sha"]), "filled disparity (matcher + temporal fill), frame %d" % t
        assert bool(g[p + "has_plane"]) == (normal is not None)
        np.testing.assert_allclose(np.asarray(normal, np.float64).reshape(3), g[p + "abc"], rtol=1e-9)
        assert st["Planar Points Before"] == int(g[p + "near_n"]) and st["Planar Points After"] == int(g[p + "road_n"])
        road = np.asarray(images["Road Image"])
        assert int((road != 0).sum()) == int(g[p + "road_px"]) and sha(road) == str(g[p + "road_sha"])
        cleaned = np.asarray(images["Filtered Road Image"])
        assert sha(cleaned) == str(g[p + "cleaned_sha"]) and int((cleaned != 0).sum()) == int(g[p + "cleaned_pts"])
        assert (st["Center Point X"], st["Center Point Y"]) == tuple(int(v) for v in g[p + "centre"])
        assert tuple(images["Normal Line"]) == tuple(int(v) for v in g[p + "normal_line"])


def test_first_frame_disparity_matches_matcher_fixture(cuda):
    """frame 0 has no previous disparity: the carried disparity is the matcher's output itself"""
    with np.load(GOLD, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    assert str(g["f0_disparity_sha"]) == str(g["f0_filled_sha"])


def test_batched_pipeline_on_the_golden_frames(cuda):
    """pipeline.StereoPipeline (the batch form: all stages for F chained frames in a handful of launches) on the same
    three frames: the disparity it carries from frame to frame is the reference orchestrator's, bit for bit; the plane
    comes from device draws, so it is checked for quality (its road mask covers the golden's road region)."""
    import np_oracle as no
    from stereo.vision_b200.masks import default_road_threshold_mask
    from stereo.vision_b200.pipeline import StereoPipeline
    with np.load(GOLD, allow_pickle=False) as z:
        g = {k: z[k] for k in z.files}
    seq, F = int(g["seq"]), int(g["n_frames"])
    frames = [synth.make_stereo_bgr(seq, t) for t in range(F)]
    L = np.stack([f[0] for f in frames])
    R = np.stack([f[1] for f in frames])
    sp = StereoPipeline(max_frames=F, device=cuda)
    out = sp.process(L, R, seed=5)
    res = out["result"]
    filled = res.filled.cpu().numpy()
    for t in range(F):
        assert sha(filled[t]) == str(g["f%d_filled_sha" % t]), "carried disparity, frame %d" % t
    assert sha(out["disparity"][0].cpu().numpy()) == str(g["f0_disparity_sha"])
    planes = res.planes.cpu().numpy()
    assert (planes[:, 7] > 0).all()
    counts = res.counts.cpu().numpy()
    tm = default_road_threshold_mask()
    for t in range(F):
        # a different winning hypothesis of the same road: the number of road points is of the golden's order
        assert 0.5 * int(g["f%d_road_n" % t]) < counts[t, 3] < 2.0 * int(g["f%d_road_n" % t])
        want, wpts = no.sanitise_road_image(res.road_mask[t].cpu().numpy(), tm)
        assert np.array_equal(out["cleaned"][t].cpu().numpy(), want)
        assert int(out["counts"][t]) == len(wpts)


def test_unmodified_orchestrator_on_gpu_helpers(cuda):
    """INTEGRATION.md section 1, executed: the reference's own stereovision.performStereoVision (byte code of the unmodified
    source, oracle/_ref/) with `functions` routed to the GPU helpers (stereo.vision_b200.dropin.route), against the same
    orchestrator on the reference's own functions module -- two chained frames, `random.seed` before each call: the tiled
    display image (every stage's picture), the carried disparity and the plane normal must agree."""
    import ref_compiled as rc
    import ref_harness as rh
    import stereo.vision_b200.dropin as dropin
    if not rc.available():
        pytest.skip("oracle/_ref was not built (python oracle/build_ref.py where /root/reference exists)")
    f_ref = rc.load_functions()
    f_ref.cv2 = rh._Cv2Compat(f_ref.cv2)              # OpenCV 4's findContours in the OpenCV 3 signature functions.py:379 expects
    f_ref2 = rc.load_functions()
    f_ref2.cv2 = rh._Cv2Compat(f_ref2.cv2)
    sv_ref = rc.load_orchestrator(f_ref)
    sv_gpu = rc.load_orchestrator(dropin.route(f_ref2))
    opts = {'crop_disparity': False, 'pause_playback': False, 'max_disparity': 128, 'ransac_trials': 600, 'road_color_thresh': 10,
            'loop': False, 'point_threshold': 0.05, 'image_tiles': True, 'img_size': (544, 1024), 'threshold_option': 'previous',
            'record_video': False, 'record_stats': False, 'video_filename': 'previous.avi'}
    o_ref, o_gpu = dict(opts), dict(opts)
    p_ref = p_gpu = None
    for t in range(2):
        imgL, imgR = synth.make_stereo_bgr(2, t)
        random.seed(synth.ransac_seed(2, t))
        tile_ref, p_ref, n_ref = sv_ref.performStereoVision(imgL.copy(), imgR.copy(), p_ref, o_ref)
        state_ref = random.getstate()
        random.seed(synth.ransac_seed(2, t))
        tile_gpu, p_gpu, n_gpu = sv_gpu.performStereoVision(imgL.copy(), imgR.copy(), p_gpu, o_gpu)
        assert random.getstate() == state_ref, "Python's generator after the frame"
        assert np.array_equal(np.asarray(p_gpu), np.asarray(p_ref)), "carried disparity, frame %d" % t
        assert n_ref is not None and n_gpu is not None
        np.testing.assert_allclose(np.asarray(n_gpu), np.asarray(n_ref), rtol=1e-9, atol=1e-13 * float(np.linalg.norm(n_ref)))
        assert tile_gpu.shape == tile_ref.shape and np.array_equal(tile_gpu, tile_ref), "tiled display image, frame %d" % t

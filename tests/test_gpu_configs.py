"""BASELINE.json's configurations at their real sizes: the 300-frame chained sequence (config 2) frame by frame
against the oracle, resident and streamed; the overlay half of a14; the 10 000 x 4 000 corner of config 5;
and the per-sequence result hash that bench.py prints per rank (config 4)."""
import hashlib

import numpy as np
import pytest

import np_oracle as no
import stereo.vision_b200.synth as synth
from helpers import check_frame, oracle_sequence, to_host

pytestmark = pytest.mark.gpu


def _carmask(shape):
    from stereo.vision_b200.masks import default_carmask
    return default_carmask(shape, prefer_cwd=False)


def _make_sequence_parallel(seq, F, shape=(544, 1024)):
    import multiprocessing as mp
    import os
    H, W = shape
    disp = np.empty((F, H, W), np.uint8)
    bgr = np.empty((F, H, W, 3), np.uint8)
    with mp.get_context("fork").Pool(min(16, os.cpu_count() or 1)) as pool:
        for t, (d, c) in enumerate(pool.starmap(synth.make_frame, [(seq, t, shape) for t in range(F)], chunksize=4)):
            disp[t], bgr[t] = d, c
    return disp, bgr


def _oracle_frame(args):
    d, c, prev_filled, cm, si, ti, T, S = args
    # the temporal fill is recomputed by the caller's chain; here the frame is judged on its own filled disparity
    o = no.run_frame(d, c, prev_filled, cm, indices=(si, ti), trials=T, n_samples=S)
    rd = o["road"]
    return {"filled": o["filled"], "N": len(o["cloud"]["d"]), "M": len(o["mcloud"]["d"]), "best": o["best"], "abc": o["abc"],
            "K": rd["near_n"], "R": rd["road_n"], "mask": np.packbits(rd["mask"]), "hist": rd["hist"]}


def test_config2_300_chained_frames_resident_and_streamed(cuda):
    """loop.py:57-78 at its real length: one 300-frame sequence (previous-disparity fill chained through all of
    it), (a) resident in one RoadPath.process call and (b) through SequenceStreamer.run in chunks of 30 (ten
    chunk boundaries), EVERY frame compared with the oracle: filled disparity, N, M, the winning hypothesis and
    plane, K, the hue histogram, R and the road mask.  Draws: the reference's own (`random.seed(ransac_seed)` per
    frame, replayed on the device; 12 frames of them are checked against CPython directly)."""
    import multiprocessing as mp
    import os
    import random
    import torch
    from stereo.vision_b200.pipeline import RoadPath, SequenceStreamer
    F, T, S, seq = 300, 600, 400, 6
    disp, bgr = _make_sequence_parallel(seq, F)
    cm = _carmask(disp.shape[1:])
    seeds = [synth.ransac_seed(seq, t) for t in range(F)]
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F)
    r = eng.process(disp, bgr, py_seeds=seeds)
    si, ti = (a.cpu().numpy() for a in r.indices)
    res = to_host(r)
    del r
    # the chain of filled disparities (a1) is sequential; everything else of a frame is independent given it
    filled = [None] * F
    prev = None
    for t in range(F):
        prev = filled[t] = no.fill_disparity(disp[t], prev)
        assert np.array_equal(res["filled"][t], prev), "filled disparity, frame %d" % t
    for t in range(0, F, 25):                                   # the replayed draws against CPython itself
        mc = no.project(no.mask_disparity(filled[t], cm))
        rng = random.Random()
        rng.seed(seeds[t])
        ws, wt = no.replay_indices(mc["xyz"], T, S, rng=rng)
        assert np.array_equal(si[t], ws) and np.array_equal(ti[t], wt), "draws of frame %d" % t
    jobs = [(disp[t], bgr[t], filled[t - 1] if t else None, cm, si[t], ti[t], T, S) for t in range(F)]
    with mp.get_context("fork").Pool(min(16, os.cpu_count() or 1)) as pool:
        orc = pool.map(_oracle_frame, jobs, chunksize=4)
    for t, o in enumerate(orc):
        N_, M_, K_, R_ = (int(v) for v in res["counts"][t])
        assert (N_, M_, K_, R_) == (o["N"], o["M"], o["K"], o["R"]), ("counts, frame %d" % t, (N_, M_, K_, R_), o)
        assert int(res["planes"][t][6]) == o["best"], "winning hypothesis, frame %d" % t
        np.testing.assert_allclose(res["planes"][t][:3], o["abc"], rtol=1e-9, atol=1e-13 * float(np.linalg.norm(o["abc"])))
        assert np.array_equal(res["hist"][t][:no.HUE_BINS].astype(np.int64), o["hist"]), "hue histogram, frame %d" % t
        assert np.array_equal(np.packbits(res["road_mask"][t]), o["mask"]), "road mask, frame %d" % t
    assert res["planes"][:, 7].all() and res["road_mask"].any()
    # (b) streamed: 10 chunks of 30, same seeds
    del eng
    torch.cuda.empty_cache()
    sr = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=30)
    out = sr.run_sync(torch.from_numpy(disp).pin_memory(), torch.from_numpy(bgr).pin_memory(), py_seeds=seeds)
    assert np.array_equal(out["counts"].numpy(), res["counts"])
    assert np.array_equal(out["planes"].numpy(), res["planes"])
    assert np.array_equal(out["road_mask"].numpy(), res["road_mask"])
    # ... and the throughput mode's draws give the same results streamed as resident (keyed by the frame number)
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F)
    dev = to_host(eng.process(disp, bgr, seed=20260000))
    out = sr.run_sync(torch.from_numpy(disp).pin_memory(), torch.from_numpy(bgr).pin_memory(), seed=20260000)
    assert np.array_equal(out["counts"].numpy(), dev["counts"])
    assert np.array_equal(out["planes"].numpy(), dev["planes"])
    assert np.array_equal(out["road_mask"].numpy(), dev["road_mask"])


@pytest.mark.parametrize("road_path", ["sites", "records"])
def test_overlay_image_road_map(cuda, road_path):
    """stereovision.py:131-133: imageRoadMap = imgL.copy(); imageRoadMap[y][x] = [0,255,0] for every road point"""
    import torch
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 3, 300, 400
    disp, bgr = synth.make_sequence(11, F)
    cm = _carmask(disp.shape[1:])
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, road_path=road_path)
    ov = torch.from_numpy(bgr).to(cuda).clone()
    r = eng.process(disp, bgr, seed=8, overlay=ov)
    si, ti = eng.dump_indices(F, seed=8)
    res = to_host(r)
    got = r.overlay.cpu().numpy()
    orc = oracle_sequence(disp, bgr, cm, indices=(si.cpu().numpy(), ti.cpu().numpy()), trials=T, n_samples=S)
    for t in range(F):
        check_frame(res, t, orc[t])
        want = bgr[t].copy()
        for x, y in orc[t]["road"]["pts"]:                      # the reference's loop, point by point
            want[y][x] = [0, 255, 0]
        assert orc[t]["road"]["road_n"] > 1000
        assert np.array_equal(got[t], want), "Image Road Map, frame %d" % t


def test_config5_corner_10000_hypotheses_4000_samples(cuda):
    """BASELINE config 5's far corner: one 2048x1088 frame, 10 000 hypotheses x 4 000 samples -- with the reference's
    draws (random.seed, replayed on the device; set path with a 16 405-entry setsize) and with device draws."""
    import random
    from stereo.vision_b200.pipeline import RoadPath
    shape, T, S = (1088, 2048), 10000, 4000
    disp, bgr = synth.make_sequence(12, 1, shape=shape)
    cm = _carmask(shape)
    eng = RoadPath(shape=shape, step=2, trials=T, samples=S, carmask=cm, max_frames=1)
    r = eng.process(disp, bgr, py_seeds=[31337])
    si, ti = (a.cpu().numpy() for a in r.indices)
    res = to_host(r)
    mc = no.project(no.mask_disparity(disp[0], cm))
    rng = random.Random()
    rng.seed(31337)
    ws, wt = no.replay_indices(mc["xyz"][:, :], 40, S, rng=rng)          # the first 40 trials against CPython (2 s)
    assert np.array_equal(si[0][:40], ws) and np.array_equal(ti[0][:40], wt)
    orc = oracle_sequence(disp, bgr, cm, indices=(si, ti), trials=T, n_samples=S)
    check_frame(res, 0, orc[0])
    r = eng.process(disp, bgr, seed=5)
    si, ti = eng.dump_indices(1, seed=5)
    res = to_host(r)
    orc = oracle_sequence(disp, bgr, cm, indices=(si.cpu().numpy(), ti.cpu().numpy()), trials=T, n_samples=S)
    check_frame(res, 0, orc[0])
    assert res["planes"][0, 7] == 1.0


def test_result_hash_is_the_single_gpu_hash(cuda):
    """a rank's results must be the single-GPU results for the same sequence id.  bench.py prints
    shard.result_sha per rank; tests/golden/result_sha.json holds the hashes of sequences 0-7 (32 frames, the
    oracle's masks, planes' winning trials and counts, made by oracle/make_result_sha.py on the CPU), so every rank of
    a multi-GPU run is compared with a value no GPU produced."""
    import json
    import os
    import stereo.vision_b200.shard as shard
    from stereo.vision_b200.pipeline import RoadPath
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "result_sha.json")
    gold = json.load(open(path))
    F, T, S = gold["frames"], gold["trials"], gold["samples"]
    cm = _carmask((544, 1024))
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F)
    for seq in (0, 3, 7):
        disp, bgr = synth.make_sequence(seq, F)
        r = eng.process(disp, bgr, py_seeds=[synth.ransac_seed(seq, t) for t in range(F)])
        assert shard.result_sha(r.road_mask, r.planes, r.counts) == gold["sha"][str(seq)], seq


def test_scores_do_not_depend_on_staging_batch_or_flags(cuda):
    """svb_ransac_score with device draws sums a hypothesis' terms in one fixed order (even raw draws into one accumulator,
    odd ones into the other): the errors are the same bits whether the cloud is staged in shared memory (max_m >= M) or
    gathered from L2 (max_m = 0: the instantiation with 16 loads in flight; max_m < M: the global path of the staged kernel),
    whether a frame is scored alone or in a batch, and with SVB_RANSAC_SHARE_SM -- which is what lets a 30-frame chunk of
    the streamer and a 300-frame resident batch be compared bit for bit."""
    import ctypes as C
    import torch
    import stereo.vision_b200._native as N
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 5, 600, 400
    disp, bgr = synth.make_sequence(9, F)
    eng = RoadPath(trials=T, samples=S, carmask=_carmask(disp.shape[1:]), max_frames=F)
    res = eng.process(disp, bgr, seed=41, frame0=3)
    want = res.errors.cpu().numpy().view(np.uint64).copy()
    m = res.counts[:, 1].cpu().numpy()
    assert (m >= S).all() and np.isfinite(res.errors.cpu().numpy()).mean() > 0.99

    def score(f0, n, max_m, flags):
        err = torch.full((n, T), -1.0, dtype=torch.float64, device=cuda)
        N.check(eng.lib.svb_ransac_score(N.ptr(eng.mcloud[f0:]), 1, eng.L, N.ptr(eng.counts[f0:]), 4, 1, n, T, S, max_m,
                                         C.byref(eng._cam), None, None, 41, 3 + f0, flags, N.ptr(err), N.stream_ptr()), "svb_ransac_score")
        torch.cuda.synchronize()
        return err.cpu().numpy().view(np.uint64)
    for max_m, flags in ((eng.max_m, 0), (0, 0), (int(m.min()) - 1, 0), (eng.max_m, N.RANSAC_SHARE_SM), (0, N.RANSAC_SHARE_SM)):
        assert np.array_equal(score(0, F, max_m, flags), want), (max_m, flags)
    for f in range(F):                                        # one frame per launch
        assert np.array_equal(score(f, 1, eng.max_m, 0), want[f:f + 1]), f


@pytest.mark.parametrize("M", [400, 401, 512, 513, 1000, 4096, 4097, 32768, 32769, 60001])
def test_device_draws_at_edge_cloud_sizes(cuda, M):
    """svb_ransac_score's own draws on a synthetic masked cloud of exactly M records: M = S (every record is drawn), a power
    of two (no draw is rejected), one more (every other draw is rejected), a cloud too large for shared memory.  The dumped
    draws are S distinct indices below M; the scores equal the oracle's (functions.py:262-276, 289) on those draws, and
    they are the same bits staged and from global memory."""
    import ctypes as C
    import torch
    import stereo.vision_b200._native as N
    from stereo.vision_b200.pipeline import RoadPath
    T, S = 96, 400
    rng = np.random.default_rng(M)
    cam = no.cam_of(None)
    eng = RoadPath(shape=(96, 160), step=2, trials=T, samples=S, carmask=np.full((96, 160), 255, np.uint8), max_frames=1)
    sites = rng.choice(1023 * 543, size=M, replace=False)
    x, y = (sites % 1023).astype(np.int64), (sites // 1023).astype(np.int64)
    inv_z = 0.02 * (x - cam.cw) / cam.f - 0.9 * (y - cam.ch) / cam.f + 0.25
    d = np.clip(np.rint(cam.f * cam.B * np.abs(inv_z)) + rng.integers(-2, 3, M), 1, 254).astype(np.int64)
    rec = torch.from_numpy((x | (y << 11) | (d << 22)).astype(np.int32)).to(cuda).reshape(1, M)
    counts = torch.tensor([[M, M, 0, 0]], dtype=torch.int32, device=cuda)
    si = torch.empty((1, T, S), dtype=torch.int32, device=cuda)
    ti = torch.empty((1, T, 3), dtype=torch.int32, device=cuda)
    st = N.stream_ptr()
    N.check(eng.lib.svb_ransac_dump_indices(N.ptr(rec), 1, M, N.ptr(counts), 4, 1, 1, T, S, C.byref(eng._cam), 77, 5,
                                            N.ptr(si), N.ptr(ti), st), "dump")
    errs = {}
    for name, max_m in (("staged", M), ("global", 0)):
        e = torch.full((1, T), -1.0, dtype=torch.float64, device=cuda)
        N.check(eng.lib.svb_ransac_score(N.ptr(rec), 1, M, N.ptr(counts), 4, 1, 1, T, S, max_m, C.byref(eng._cam), None, None,
                                         77, 5, 0, N.ptr(e), st), "score")
        errs[name] = e.cpu().numpy()[0]
    ex = torch.full((1, T), -1.0, dtype=torch.float64, device=cuda)             # the warp-per-hypothesis scorer on the same draws
    N.check(eng.lib.svb_ransac_score(N.ptr(rec), 1, M, N.ptr(counts), 4, 1, 1, T, S, M, C.byref(eng._cam), N.ptr(si), N.ptr(ti),
                                     0, 0, 0, N.ptr(ex), st), "score explicit")
    torch.cuda.synchronize()
    s, t3 = si.cpu().numpy()[0], ti.cpu().numpy()[0]
    assert s.min() >= 0 and s.max() < M and (np.diff(np.sort(s, axis=1), axis=1) > 0).all()
    if M == S:
        assert (np.sort(s, axis=1) == np.arange(M)).all()
    assert np.array_equal(errs["staged"].view(np.uint64), errs["global"].view(np.uint64))
    Z = cam.f * cam.B / d.astype(np.float64)
    xyz = np.stack([((x - cam.cw) * Z) / cam.f, ((y - cam.ch) * Z) / cam.f, Z], axis=1)
    _, _, want = no.ransac_from_indices(xyz, s, t3, sites=np.stack([y, x], axis=1))
    got = errs["staged"]
    assert np.array_equal(np.isinf(got), np.isinf(want)) and np.isfinite(got).sum() > T // 2
    fin = np.isfinite(want)
    assert np.allclose(got[fin], want[fin], rtol=1e-9, atol=0) and np.allclose(ex.cpu().numpy()[0][fin], want[fin], rtol=1e-9, atol=0)

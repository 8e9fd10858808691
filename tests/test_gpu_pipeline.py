"""GPU parity of the fused path (RoadPath.process) against the numpy oracle."""
import numpy as np
import pytest

import np_oracle as no
import stereo.vision_b200.synth as synth
from helpers import check_frame, oracle_sequence, to_host

pytestmark = pytest.mark.gpu


def _carmask(shape):
    from stereo.vision_b200.masks import default_carmask
    return default_carmask(shape, prefer_cwd=False)


def test_sequence_python_replay_indices(cuda):
    """3 chained frames, full size, draws = replay of Python's random under the frame seeds."""
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 3, 600, 600
    disp, bgr = synth.make_sequence(0, F)
    cm = _carmask(disp.shape[1:])
    seeds = [synth.ransac_seed(0, t) for t in range(F)]
    orc = oracle_sequence(disp, bgr, cm, seeds=seeds, trials=T, n_samples=S)
    si = np.stack([o["indices"][0] for o in orc])
    ti = np.stack([o["indices"][1] for o in orc])
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F)
    res = to_host(eng.process(disp, bgr, indices=(si, ti)))
    for t in range(F):
        check_frame(res, t, orc[t])


@pytest.mark.parametrize("S", [400, 600])
def test_sequence_device_draws(cuda, S):
    """Counter-based on-device draws: dump them, feed the oracle the same indices, compare."""
    from stereo.vision_b200.pipeline import RoadPath
    F, T = 4, 600
    disp, bgr = synth.make_sequence(1, F)
    cm = _carmask(disp.shape[1:])
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F)
    r = eng.process(disp, bgr, seed=1234, frame0=7)
    si, ti = eng.dump_indices(F, seed=1234, frame0=7)
    si, ti = si.cpu().numpy(), ti.cpu().numpy()
    res = to_host(r)
    M = res["counts"][:, 1]
    for t in range(F):
        assert si[t].min() >= 0 and si[t].max() < M[t]
        srt = np.sort(si[t], axis=1)
        assert (np.diff(srt, axis=1) > 0).all(), "samples of a trial must be distinct"
    orc = oracle_sequence(disp, bgr, cm, indices=(si, ti), trials=T, n_samples=S)
    for t in range(F):
        check_frame(res, t, orc[t])


@pytest.mark.parametrize("shape,step", [((37, 53), 1), ((37, 53), 2), ((64, 96), 3), ((48, 80), 2), ((33, 64), 1)])
def test_small_ragged_shapes(cuda, shape, step):
    """Unaligned widths (scalar loads), odd steps, tiny clouds; draws from the device."""
    from stereo.vision_b200.pipeline import RoadPath
    rng = np.random.default_rng(5)
    F, T, S = 3, 64, 40
    H, W = shape
    disp = rng.integers(0, 60, size=(F, H, W), dtype=np.uint8)
    disp[rng.random((F, H, W)) < 0.3] = 0
    bgr = rng.integers(0, 256, size=(F, H, W, 3), dtype=np.uint8)
    cm = (rng.random((H, W)) < 0.7).astype(np.uint8) * 200
    eng = RoadPath(shape=shape, step=step, trials=T, samples=S, carmask=cm, max_frames=F, point_threshold=0.5,
                   colour_threshold=0)
    r = eng.process(disp, bgr, seed=9)
    si, ti = eng.dump_indices(F, seed=9)
    res = to_host(r)
    orc = oracle_sequence(disp, bgr, cm, indices=(si.cpu().numpy(), ti.cpu().numpy()), trials=T, n_samples=S,
                          step=step, threshold=0.5, colour_threshold=0)
    for t in range(F):
        check_frame(res, t, orc[t])
    # the streamed road path (used for lattices too large for the cluster kernel) on the same ragged input
    import stereo.vision_b200._native as N
    N.load().svb_debug_force_two_pass(3)
    try:
        res3 = to_host(eng.process(disp, bgr, seed=9))
    finally:
        N.load().svb_debug_force_two_pass(0)
    for k in ("road_mask", "counts", "hist"):
        assert np.array_equal(res[k], res3[k]), k


def test_two_half_batches_on_two_streams_equal_one_batch(cuda):
    """RoadPath(overlap=True), the default, runs a batch of >= 192 frames (here: OVERLAP_MIN_FRAMES lowered) as two half-batches on two streams (the scorer
    of one half beside the other half's kernels): every output equals the one-batch run's, for device draws, the reference's draws
    (per-frame seeds) and explicit indices; an odd frame count; two calls in a row (buffers of the last call in flight)."""
    import torch
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 37, 300, 400
    disp, bgr = synth.make_sequence(14, F)
    cm = _carmask(disp.shape[1:])
    one = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, overlap=False)
    two = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, overlap=True)
    two.OVERLAP_MIN_FRAMES = 16
    seeds = [synth.ransac_seed(14, t) for t in range(F)]
    for kw in (dict(seed=77, frame0=5), dict(py_seeds=seeds), dict(seed=78)):
        a = one.process(disp, bgr, **kw)
        b = two.process(disp, bgr, **kw)
        assert two._side is not None
        for k in ("filled", "counts", "planes", "errors", "hist", "road_mask", "cloud", "mcloud"):
            ta, tb = getattr(a, k), getattr(b, k)
            if k in ("cloud", "mcloud"):                    # slabs: only the first N / M records of a frame are defined
                n = a.counts[:, 0 if k == "cloud" else 1].cpu().numpy()
                for t in range(F):
                    assert torch.equal(ta[t, :n[t]], tb[t, :n[t]]), (k, t)
            else:
                assert torch.equal(ta, tb), (k, list(kw))
        if "py_seeds" in kw:
            assert torch.equal(a.indices[0], b.indices[0]) and torch.equal(a.indices[1], b.indices[1])
            assert torch.equal(one.mt_state[:F], two.mt_state[:F])
    si, ti = one.dump_indices(F, seed=78)
    c = two.process(disp, bgr, indices=(si, ti))
    # explicit indices are summed in another order than device draws: the winner's error agrees to rounding, the rest exactly
    assert torch.equal(c.planes[:, :5], a.planes[:, :5]) and torch.equal(c.planes[:, 6:], a.planes[:, 6:])
    assert torch.allclose(c.planes[:, 5], a.planes[:, 5], rtol=1e-12, atol=0) and torch.equal(c.road_mask, a.road_mask)
    assert (a.planes[:, 7] == 1).all() and a.road_mask.any()


def test_chained_calls_on_one_engine_with_its_own_filled_frame(cuda):
    """prev = result.filled[-1] of the SAME engine is a view of the buffer the next fill overwrites (process() clones it):
    two chained calls equal one call over all frames"""
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 6, 200, 400
    disp, bgr = synth.make_sequence(16, F)
    cm = _carmask(disp.shape[1:])
    whole = to_host(RoadPath(trials=T, samples=S, carmask=cm, max_frames=F).process(disp, bgr, seed=9))
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=3)
    a = eng.process(disp[:3], bgr[:3], seed=9)
    first = to_host(a)
    b = to_host(eng.process(disp[3:], bgr[3:], prev=a.filled[-1], seed=9, frame0=3))
    for k in ("filled", "counts", "planes", "hist", "road_mask"):
        assert np.array_equal(np.concatenate([first[k], b[k]]), whole[k]), k


def test_repeated_runs_are_bit_identical(cuda):
    """compute-sanitizer (racecheck) is closed on the GPU pool; what a data race in the compaction, the shared-memory
    histograms, the staged scorer or the scatter would do is change a result from run to run.  Ten runs of one batch on two
    engines (fresh and reused buffers), both road paths: every output identical, every time."""
    import torch
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 12, 600, 400
    disp, bgr = synth.make_sequence(15, F)
    cm = _carmask(disp.shape[1:])
    ov0 = torch.from_numpy(bgr).to(cuda)
    for road_path in ("sites", "records"):
        first = None
        for run in range(10):
            eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, road_path=road_path) if run % 5 == 0 else eng
            ov = ov0.clone()
            r = eng.process(disp, bgr, seed=3, overlay=ov)
            n = r.counts[:, :2].cpu().numpy()
            got = {k: getattr(r, k).clone() for k in ("filled", "counts", "planes", "errors", "hist", "road_mask")}
            got["overlay"] = ov
            got["cloud"] = torch.cat([r.cloud[t, :n[t, 0]] for t in range(F)])
            got["mcloud"] = torch.cat([r.mcloud[t, :n[t, 1]] for t in range(F)])
            if first is None:
                first = got
            else:
                for k, v in got.items():
                    assert torch.equal(v, first[k]), (road_path, run, k)
        assert (first["planes"][:, 7] == 1).all() and first["road_mask"].any()


def test_no_plane_when_masked_cloud_too_small(cuda):
    """M < S: every trial raises in the reference -> RANSAC returns (None, None) -> black mask."""
    from stereo.vision_b200.pipeline import RoadPath
    disp, bgr = synth.make_sequence(2, 1)
    cm = np.zeros(disp.shape[1:], np.uint8)
    cm[200:210, 100:160] = 255                       # 5 x 30 lattice sites at most
    eng = RoadPath(carmask=cm, max_frames=1)
    res = to_host(eng.process(disp, bgr, seed=3))
    orc = oracle_sequence(disp, bgr, cm, seeds=[0])
    assert orc[0]["abc"] is None
    check_frame(res, 0, orc[0])


def test_hue_bins_exhaustive(cuda):
    """a10 known-answer test over ALL 2^24 colours: kernel hue bin == oracle hue bin."""
    import ctypes as C
    import torch
    import stereo.vision_b200._native as N
    lib = N.load()
    for r0 in range(0, 256, 32):                                  # 32 red values (2M colours) per launch
        r, g, b = np.meshgrid(np.arange(r0, r0 + 32, dtype=np.uint8), np.arange(256, dtype=np.uint8),
                              np.arange(256, dtype=np.uint8), indexing="ij")
        rgb = np.stack([r.ravel(), g.ravel(), b.ravel()], axis=1)
        t = torch.from_numpy(rgb).to(cuda)
        out = torch.empty((len(rgb),), dtype=torch.int16, device=cuda)
        N.check(lib.svb_hue_bins(N.ptr(t), len(rgb), N.ptr(out), N.stream_ptr()), "svb_hue_bins")
        got = out.cpu().numpy().astype(np.int32)
        want = no.hue_bin(rgb[:, 0], rgb[:, 1], rgb[:, 2])
        bad = np.nonzero(got != want)[0]
        assert len(bad) == 0, (len(bad), rgb[bad[:5]], got[bad[:5]], want[bad[:5]])


def test_reprojection_exhaustive(cuda):
    """a13 known-answer test over every (x,d) and (y,d) pair: the kernel's float64
    sequence and the displacement-bit table both equal the oracle's reprojection."""
    import ctypes as C
    import torch
    import stereo.vision_b200._native as N
    from stereo.vision_b200.camera import DEFAULT_CAMERA
    lib, cam = N.load(), N.ccam(DEFAULT_CAMERA)
    H, W = 544, 1024
    bits = torch.empty((lib.svb_reproject_bits_words(H, W),), dtype=torch.int32, device=cuda)
    ok = torch.zeros((1,), dtype=torch.int32, device=cuda)
    N.check(lib.svb_reproject_bits(H, W, C.byref(cam), N.ptr(bits), N.ptr(ok), N.stream_ptr()), "bits")
    assert int(ok.item()) == 1
    bits = bits.cpu().numpy().astype(np.uint32)
    c = no.cam_of(None)
    for axis, n, centre, row0 in ((0, W, c.cw, 0), (1, H, c.ch, W)):
        out = torch.empty((n, 256), dtype=torch.int32, device=cuda)
        N.check(lib.svb_reproject_table(axis, n, C.byref(cam), N.ptr(out), N.stream_ptr()), "table")
        got = out.cpu().numpy()[:, 1:]
        coord = np.arange(n, dtype=np.float64)[:, None]
        d = np.arange(1, 256, dtype=np.float64)[None, :]
        Z = (c.f * c.B) / d
        V = ((coord - centre) * Z) / c.f
        want = (((V * c.f) / Z) + centre).astype(np.int32)
        assert np.array_equal(got, want)
        delta = np.arange(n)[:, None] - want
        assert set(np.unique(delta)) <= {0, 1} and (delta == 1).sum() > 0
        di = np.arange(1, 256)
        tb = (bits[(row0 + np.arange(n))[:, None] * 8 + (di >> 5)[None, :]] >> (di & 31)[None, :].astype(np.uint32)) & 1
        assert np.array_equal(tb.astype(np.int64), delta)


def test_road_points_without_reprojection_table(cuda):
    """The per-point float64 reprojection path (no table) gives the same masks."""
    from stereo.vision_b200.pipeline import RoadPath
    F = 2
    disp, bgr = synth.make_sequence(4, F)
    cm = _carmask(disp.shape[1:])
    a = RoadPath(carmask=cm, max_frames=F, samples=400)
    b = RoadPath(carmask=cm, max_frames=F, samples=400, use_reproj_table=False)
    assert a.reproj_bits is not None and b.reproj_bits is None
    ra, rb = to_host(a.process(disp, bgr, seed=5)), to_host(b.process(disp, bgr, seed=5))
    assert np.array_equal(ra["road_mask"], rb["road_mask"]) and np.array_equal(ra["counts"], rb["counts"])
    assert ra["road_mask"].any()


def test_road_paths_agree(cuda):
    """svb_road_points' three implementations -- cluster kernel, legacy two-pass, streamed -- and
    svb_road_points_sites (the default of RoadPath) agree."""
    import stereo.vision_b200._native as N
    from stereo.vision_b200.pipeline import RoadPath
    F = 3
    disp, bgr = synth.make_sequence(5, F)
    cm = _carmask(disp.shape[1:])
    eng = RoadPath(carmask=cm, max_frames=F, samples=400, road_path="records")
    assert eng.fused_road and not eng.sites_road
    ra = to_host(eng.process(disp, bgr, seed=5))
    sites = RoadPath(carmask=cm, max_frames=F, samples=400)           # the default: a8-a14 from disparity + colour
    assert sites.sites_road
    rs = to_host(sites.process(disp, bgr, seed=5))
    for k in ("road_mask", "counts", "planes"):
        assert np.array_equal(ra[k], rs[k]), ("sites", k)
    assert np.array_equal(ra["hist"][:, :no.HUE_BINS], rs["hist"][:, :no.HUE_BINS])
    # slot SVB_HIST_SPANS of a row: the 16-pixel colour spans the sites kernel read (at least one per 8 near points at step 2)
    spans = rs["hist"][:, 1023]
    assert (ra["hist"][:, no.HUE_BINS:] == 0).all() and (rs["hist"][:, no.HUE_BINS:1023] == 0).all()
    assert (spans * 8 >= rs["counts"][:, 2]).all() and (spans <= 272 * 64).all() and (spans > 0).all()
    for mode in (1, 3):
        N.load().svb_debug_force_two_pass(mode)
        try:
            rb = to_host(eng.process(disp, bgr, seed=5))
        finally:
            N.load().svb_debug_force_two_pass(0)
        for k in ("road_mask", "counts", "hist"):
            assert np.array_equal(ra[k], rb[k]), (mode, k)
    assert ra["road_mask"].any()


@pytest.mark.parametrize("road_path", ["sites", "records"])
def test_config3_full_density_step1(cuda, road_path):
    """BASELINE config 3: step 1 (555 489 lattice sites), RGB attach, masked-cloud RANSAC.  The road points
    come from the disparity + colour image (default), or from the cloud records on the streamed path (the near
    list of a frame no longer fits a cluster's shared memory)."""
    from stereo.vision_b200.pipeline import RoadPath
    F, T, S = 2, 600, 400
    disp, bgr = synth.make_sequence(8, F)
    cm = _carmask(disp.shape[1:])
    eng = RoadPath(step=1, trials=T, samples=S, carmask=cm, max_frames=F, road_path=road_path)
    assert not eng.fused_road and eng.road_launches == (2 if road_path == "sites" else 3)
    assert eng.sites_road == (road_path == "sites")
    r = eng.process(disp, bgr, seed=31)
    si, ti = eng.dump_indices(F, seed=31)
    res = to_host(r)
    orc = oracle_sequence(disp, bgr, cm, indices=(si.cpu().numpy(), ti.cpu().numpy()), trials=T, n_samples=S, step=1)
    for t in range(F):
        check_frame(res, t, orc[t])


def test_config5_stress_2048x1088(cuda):
    """BASELINE config 5: 2048x1088 frames (masks scaled x2), larger RANSAC (2000 hypotheses x 1000 samples)."""
    from stereo.vision_b200.pipeline import RoadPath
    shape, F, T, S = (1088, 2048), 2, 2000, 1000
    disp, bgr = synth.make_sequence(9, F, shape=shape)
    cm = _carmask(shape)
    eng = RoadPath(shape=shape, step=2, trials=T, samples=S, carmask=cm, max_frames=F)
    r = eng.process(disp, bgr, seed=77)
    si, ti = eng.dump_indices(F, seed=77)
    res = to_host(r)
    orc = oracle_sequence(disp, bgr, cm, indices=(si.cpu().numpy(), ti.cpu().numpy()), trials=T, n_samples=S, step=2)
    for t in range(F):
        check_frame(res, t, orc[t])
    assert res["planes"][:, 7].all()


def test_sequence_streamer_matches_resident_path(cuda):
    """The host-to-host entry point (chunked, three streams, lattice-row upload of the colour image) gives the
    results of one resident RoadPath.process call over the same sequence."""
    import torch
    from stereo.vision_b200.pipeline import RoadPath, SequenceStreamer
    F, T, S = 7, 200, 300
    disp, bgr = synth.make_sequence(4, F)
    cm = _carmask(disp.shape[1:])
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F)
    ref = to_host(eng.process(disp, bgr, seed=21))
    sr = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=3, upload="full")
    out = sr.run_sync(torch.from_numpy(disp).pin_memory(), torch.from_numpy(bgr).pin_memory(), seed=21)
    assert np.array_equal(out["counts"].numpy(), ref["counts"])
    assert np.array_equal(out["road_mask"].numpy(), ref["road_mask"])
    assert np.array_equal(out["planes"].numpy(), ref["planes"])
    assert sr.h2d_bytes == F * (544 * 1024 + 272 * 1024 * 3) and ref["road_mask"].any()
    # upload="lattice": only the lattice rows of the disparity travel as well.  N (cloud points) is still the full count:
    # it is a count over lattice sites
    sl = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=3, upload="lattice")
    outl = sl.run_sync(torch.from_numpy(disp).pin_memory(), torch.from_numpy(bgr).pin_memory(), seed=21)
    for k in ("counts", "road_mask", "planes"):
        assert np.array_equal(outl[k].numpy(), ref[k]), k
    assert sl.h2d_bytes == F * (272 * 1024 + 272 * 1024 * 3)
    # the engines' filled frames are right on the lattice rows: engine 1 still holds frames 3-5, engine 0 frame 6
    for t, k, j in ((3, 1, 0), (4, 1, 1), (5, 1, 2), (6, 0, 0)):
        assert np.array_equal(sl.engines[k].filled[j, ::2].cpu().numpy(), ref["filled"][t, ::2]), t
    # bit-packed masks over the link, and an engine that emits no cloud (the streamer's default) against one that does
    assert not sl.engines[0].emit_cloud and sl.engines[0].cloud.numel() == 0 and eng.emit_cloud
    sp = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=3, pack_mask=True, upload="lattice")
    outp = sp.run_sync(torch.from_numpy(disp).pin_memory(), torch.from_numpy(bgr).pin_memory(), seed=21)
    assert "road_mask" not in outp and outp["road_bits"].shape == (F, 544 * 32)
    assert np.array_equal(sp.unpack_road_mask(outp["road_bits"]), ref["road_mask"])
    assert np.array_equal(outp["counts"].numpy(), ref["counts"]) and np.array_equal(outp["planes"].numpy(), ref["planes"])
    assert sp.d2h_bytes == F * (544 * 32 * 4 + 8 * 8 + 4 * 4)
    nocloud = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, emit_cloud=False)
    rn = nocloud.process(disp, bgr, seed=21)
    assert rn.cloud is None
    for k in ("counts", "planes", "errors", "hist", "road_mask", "filled"):
        assert np.array_equal(getattr(rn, k).cpu().numpy(), ref[k]), k
    with pytest.raises(ValueError):
        RoadPath(carmask=cm, max_frames=1, emit_cloud=False, road_path="records")
    out2 = sr.run_sync(disp, bgr, seed=21, out=out)               # pageable numpy input, buffers reused
    assert np.array_equal(out2["road_mask"].numpy(), ref["road_mask"])


def test_road_sites_writes_stay_inside_their_buffers(cuda):
    """svb_road_points_sites through the C-ABI with sentinel margins around every output and the scratch
    (no compute-sanitizer on the GPU pool), for the reference size and a ragged one; results equal RoadPath's."""
    import ctypes as C
    import torch
    import stereo.vision_b200._native as N
    from helpers import RedZone
    from stereo.vision_b200.pipeline import RoadPath
    lib = N.load()
    for shape, step in (((544, 1024), 2), ((75, 130), 1), ((64, 1040), 2)):
        H, W = shape
        F = 2
        if shape == (544, 1024):
            disp, bgr = synth.make_sequence(9, F)
            cm = _carmask(shape)
        else:
            rng = np.random.default_rng(H)
            disp = rng.integers(20, 60, size=(F, H, W), dtype=np.uint8)
            disp[rng.random((F, H, W)) < 0.2] = 0
            bgr = rng.integers(0, 256, size=(F, H, W, 3), dtype=np.uint8)
            cm = np.full((H, W), 255, np.uint8)
        eng = RoadPath(shape=shape, step=step, trials=64, samples=40, carmask=cm, max_frames=F, point_threshold=0.3, colour_threshold=1)
        assert eng.sites_road
        r = eng.process(disp, bgr, seed=3)
        torch.cuda.synchronize()
        rz = RedZone(cuda)
        hist = rz.alloc((F, 1024), torch.int32, -1)
        mask = rz.alloc((F, H, W), torch.uint8, 9)
        counts = rz.alloc((F, 4), torch.int32, 0)
        nb = lib.svb_road_sites_scratch_bytes(F, H, W, step)
        scratch = rz.alloc((nb,), torch.uint8)
        N.check(lib.svb_road_points_sites(N.ptr(eng.filled), N.ptr(torch.from_numpy(bgr).to(cuda)), N.ptr(eng.planes), F, H, W, step,
                                          C.byref(eng._cam), eng.colour_threshold, N.ptr(eng.reproj_bits), N.ptr(hist), N.ptr(mask), 0,
                                          N.ptr(counts), N.ptr(scratch), nb, 0), "svb_road_points_sites")
        torch.cuda.synchronize()
        rz.check()
        assert torch.equal(mask, r.road_mask) and torch.equal(hist, r.hist[:F]) and torch.equal(counts[:, 2:], r.counts[:, 2:])


def test_streamer_sends_only_the_colour_rows_the_planes_allow(cuda):
    """SequenceStreamer(upload="needed") and "pull" (the default): disparity rows up, a1-a7, planes down, near_row_ranges(), only the
    colour rows that can hold a near-plane point up, a8-a14.  Same masks, planes and counts as upload="lattice" -- with
    device draws and with the reference's, for chunks that do not divide the sequence, a first frame without a plane
    (nothing to upload for it), runs back to back, and with the colour buffers full of garbage outside the uploaded rows."""
    import torch
    from stereo.vision_b200.pipeline import RoadPath, SequenceStreamer
    F, T, S = 11, 200, 300
    disp, bgr = synth.make_sequence(5, F)
    disp[0] = 0                                                     # prev = None and no disparity: frame 0 has no cloud, no plane
    cm = _carmask(disp.shape[1:])
    dp, bp = torch.from_numpy(disp).pin_memory(), torch.from_numpy(bgr).pin_memory()
    seeds = [synth.ransac_seed(5, t) for t in range(F)]
    for kw in (dict(seed=3), dict(py_seeds=seeds)):
        ref = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=4, upload="lattice").run_sync(dp, bp, **kw)
        assert ref["planes"][0, 7] == 0 and (ref["planes"][1:, 7] == 1).all() and ref["road_mask"][1:].any()
        for chunk, mode in ((3, "needed"), (4, "needed"), (16, "needed"), (3, "pull"), (4, "auto"), (16, "pull")):
            sn = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=chunk, upload=mode)
            assert sn.upload == ("pull" if mode == "auto" else mode)        # "pull": the road kernel reads the pinned frames in place
            for c in sn.cin:
                c.random_(0, 256)                                    # whatever is not uploaded must not matter
            out = None
            for _ in range(2):
                out = sn.run_sync(dp, bp, out=out, **kw)
                for k in ("counts", "road_mask"):
                    assert torch.equal(out[k], ref[k]), (k, chunk, list(kw))
                assert torch.equal(out["planes"].view(torch.int64), ref["planes"].view(torch.int64)), (chunk, list(kw))   # bits: frame 0 is NaN
            lattice = F * (272 * 1024 + 272 * 1024 * 3)
            assert F * 272 * 1024 < sn.h2d_bytes < (0.8 if sn.upload == "needed" else 0.7) * lattice, (sn.upload, sn.h2d_bytes, lattice)
            if sn.upload == "pull":                                  # 48 bytes per span the kernel reports, at least one per 8 near points
                assert (sn.h2d_bytes - F * 272 * 1024) % 48 == 0 and sn.h2d_bytes - F * 272 * 1024 >= 48 * int(ref["counts"][:, 2].sum()) // 8
    # bit-packed masks, pageable numpy input (cannot be read in place: the colour rows are copied as with upload="needed")
    sp = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=5, pack_mask=True)
    outp = sp.run_sync(disp, bgr, seed=3)
    assert sp.upload == "pull" and sp.h2d_bytes > F * 272 * 1024
    ref = SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=5, upload="lattice").run_sync(dp, bp, seed=3)
    assert np.array_equal(sp.unpack_road_mask(outp["road_bits"]), ref["road_mask"].numpy())
    # the two halves of RoadPath.process on their own
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, emit_cloud=False)
    r = eng.process(disp, bgr, seed=3)
    whole = {k: getattr(r, k).cpu().numpy().copy() for k in ("counts", "planes", "road_mask", "hist")}
    half = eng.process(disp, None, seed=3, road=False)
    assert np.array_equal(half.planes.cpu().numpy().view(np.int64), whole["planes"].view(np.int64))
    eng.process_road(torch.from_numpy(bgr).to(cuda))
    for k in ("counts", "planes", "road_mask", "hist"):
        assert np.array_equal(getattr(half, k).cpu().numpy().view(np.uint8), whole[k].view(np.uint8)), k
    with pytest.raises(RuntimeError):
        eng.process_road(torch.from_numpy(bgr).to(cuda))
    with pytest.raises(ValueError):
        RoadPath(trials=T, samples=S, carmask=cm, max_frames=F).process(disp, None, seed=3, road=False)
    with pytest.raises(ValueError):
        SequenceStreamer(trials=T, samples=S, carmask=cm, chunk=4, upload="needed", step=1)

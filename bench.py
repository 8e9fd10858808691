#!/usr/bin/env python
"""Benchmark of the hot path (project + RANSAC + road points) on synthetic 1024x544 sequences.

    python bench.py --gpus N --steps K --warmup W            # this repo's arm (B200 kernels)
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port on host cores

A "step" is one pass of the path over one resident sequence of --frames frames (BASELINE config 2:
300 chained frames, previous-disparity fill, step 2, 600 trials x 400 samples).  `value` is frames/s
with the inputs already in HBM (they are 668 MB per sequence, larger than the 126 MB L2, so every
step streams them from DRAM again); `e2e` is the same metric through RoadPath.process with pinned
HOST buffers, host->device copies of the frames and device->host reads of masks/planes/counts inside
the timed region.  One rank per GPU, one independent sequence per rank, no hot-path collective.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W = 544, 1024
METRIC = "frames/sec @1024x544 (project+RANSAC+road pts)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--trials", type=int, default=600)
    ap.add_argument("--samples", type=int, default=400)
    ap.add_argument("--step", type=int, default=2)
    ap.add_argument("--chunk", type=int, default=60, help="frames per host->device chunk in the e2e leg (30: 65.1 k frames/s, 60: 66.3 k, 100: 66.7 k)")
    ap.add_argument("--cpu-frames", type=int, default=0, help="frames of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--path-only", action="store_true", help="skip the next_rows legs (cleaning, front end, matcher, whole frame)")
    return ap.parse_args()


# ------------------------------------------------------------------ synthetic input
def _gen_one(args):
    import stereo.vision_b200.synth as synth
    seq, t = args
    return synth.make_frame(seq, t, (H, W))


def make_sequence_parallel(seq, n_frames, workers):
    """disp [F,H,W], bgr [F,H,W,3] of sequence `seq`, generated on `workers` processes."""
    import multiprocessing as mp
    disp = np.empty((n_frames, H, W), np.uint8)
    bgr = np.empty((n_frames, H, W, 3), np.uint8)
    if workers <= 1:
        for t in range(n_frames):
            disp[t], bgr[t] = _gen_one((seq, t))
        return disp, bgr
    with mp.get_context("fork").Pool(workers) as pool:
        for t, (d, c) in enumerate(pool.imap(_gen_one, [(seq, t) for t in range(n_frames)], chunksize=4)):
            disp[t], bgr[t] = d, c
    return disp, bgr


# ------------------------------------------------------------------ CPU arm: the reference itself, else the oracle port
def _reference_kind(samples, step):
    """ "reference": oracle/_ref holds the reference's own byte code (oracle/build_ref.py) for this configuration;
    "port": the numpy restatement (oracle/np_oracle.py)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    try:
        import ref_compiled as rc
        if rc.available() and samples in (400, 600) and step == 2:
            return "reference"
    except Exception:
        pass
    return "port"


def _reference_chain(job):
    """`n` chained frames of sequence `seq` through the REFERENCE's helpers in the orchestrator's order
    (stereovision.py:58-140; functions.py compiled to byte code by oracle/build_ref.py); returns seconds spent."""
    seq, n, trials, samples, step = job
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import random
    import warnings
    import ref_compiled as rc
    import stereo.vision_b200.synth as synth
    f = rc.load_functions(samples)
    frames = [synth.make_frame(seq, t, (H, W)) for t in range(n)]
    prev = None
    t0 = time.perf_counter()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for t, (disparity, imgL) in enumerate(frames):
            random.seed(synth.ransac_seed(seq, t))
            disparity = f.disparityCleaning(disparity.copy(), 'previous', prev)          # :58
            prev = disparity                                                              # :60
            maskedDisparity = f.maskDisparity(disparity)                                  # :74
            cappedDisparity = f.capDisparity(disparity)                                   # :76
            points = f.projectDisparityTo3d(cappedDisparity, 128, imgL)                   # :84
            maskpoints = f.projectDisparityTo3d(maskedDisparity, 128)                     # :85
            planePoints = []
            try:
                normal, abc = f.RANSAC(maskpoints, trials)                                # :94
                pointDifferences = f.calculatePointErrors(abc, points)                    # :97
                points = f.computePlanarThreshold(points, pointDifferences, 0.05)         # :100
                histogram = f.calculateColourHistogram(points)                            # :103
                points = f.filterPointsByHistogram(points, histogram, 10)                 # :106
                planePoints = f.project3DPointsTo2DImagePoints(points)                    # :111
                planePoints = np.array(planePoints, np.int32).reshape((-1, 1, 2))         # :112-113
            except Exception:
                pass
            f.generatePointsAsImage(planePoints)                                          # :140
    return time.perf_counter() - t0


def _oracle_chain(job):
    """Run `n` chained frames of sequence `seq` through the numpy oracle; returns seconds spent."""
    seq, n, trials, samples, step = job
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import np_oracle as no
    import stereo.vision_b200.synth as synth
    from stereo.vision_b200.masks import default_carmask
    cm = default_carmask((H, W), prefer_cwd=False)
    frames = [synth.make_frame(seq, t, (H, W)) for t in range(n)]
    prev = None
    t0 = time.perf_counter()
    for t, (d, c) in enumerate(frames):
        o = no.run_frame(d, c, prev, cm, seed=synth.ransac_seed(seq, t), trials=trials, n_samples=samples, step=step)
        prev = o["filled"]
    return time.perf_counter() - t0


def cpu_throughput(kind, frames_per_proc, procs, trials, samples, step):
    """frames/s of the CPU arm with one independent sequence per process."""
    import multiprocessing as mp
    fn = _reference_chain if kind == "reference" else _oracle_chain
    jobs = [(1000 + p, frames_per_proc, trials, samples, step) for p in range(procs)]
    if procs == 1:
        times = [fn(jobs[0])]
    else:
        with mp.get_context("fork").Pool(procs) as pool:
            times = pool.map(fn, jobs)
    # every worker times only its helper calls (frame synthesis excluded); the job ends with the slowest
    wall = max(times)
    return frames_per_proc * procs / wall, wall


def cpu_sample_text(kind, procs, fpp, wall):
    what = ("the reference's own helpers (functions.py compiled to byte code under oracle/_ref/, stereovision.py:58-140's order, "
            "random.seed per frame)" if kind == "reference" else "oracle/np_oracle.py (numpy float64 restatement)")
    return "%d independent sequences x %d chained frames through %s, one process per core, %.1f s wall" % (procs, fpp, what, wall)


# ------------------------------------------------------------------ clocks during the timed region
class ClockSampler:
    def __init__(self, index):
        self.index, self.samples, self.reasons, self._stop, self._thr = index, [], set(), threading.Event(), None
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake_slowdown": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.002)

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------ this repo's arm
def algorithmic_bytes(kernel, L, N, M, K, R, T, S):
    """Algorithmic bytes per FRAME of each kernel (DESIGN.md "Kernels"; SURVEY 8d restated for the
    8-byte cloud / 4-byte mcloud records).  N, M, K, R are per-frame means."""
    table = {
        "fill_seq": 2.0 * H * W,                              # raw in + filled out (prev lives in registers); also counts sites
        "proj_count": 2.0 * L,                                # d + carmask at lattice sites (stand-alone projection only)
        "proj_write": 2.0 * L + 3.0 * N + 8.0 * N + 4.0 * M,  # d, carmask, rgb in; cloud + mcloud out
        "ransac_score": 4.0 * T * (S + 3) + 8.0 * T,          # gathered records (shared memory / L2) + errors out
        "road_cluster": 8.0 * N + 1.0 * H * W + 1.0 * R,      # records in, mask cleared, road bytes out
        "road_sites": 1.0 * L + 3.0 * K + 1.0 * H * W + 4.0 * K,  # lattice disparities + colour of the near points in, mask cleared, near list out
        "road_sites_scatter": 4.0 * K + 1.0 * R,              # listed near points in, road bytes out
        "road_near": 8.0 * N + 1.0 * H * W + 4.0 * K,         # records in, mask cleared, near list out
        "road_scatter": 4.0 * K + 1.0 * R,                    # near list in, road bytes out
        "road_pass1": 8.0 * N + N / 8.0,                      # records in, near bits out
        "road_pass2": N / 8.0 + 8.0 * K + 1.0 * R,            # near bits, near records in, mask bytes out
        "road_clear": 1.0 * H * W,                            # mask clear (memset node)
    }
    return table.get(kernel)


def run_b200(args):
    import torch
    import stereo.vision_b200.shard as shard
    rank, world, local = shard.rank_world()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    shard.init("nccl", dev)
    import stereo.vision_b200._native as N
    from stereo.vision_b200.pipeline import RoadPath, KERNELS_PER_BATCH
    from stereo.vision_b200.masks import default_carmask

    F, T, S = args.frames, args.trials, args.samples
    ncpu = os.cpu_count() or 1
    disp_h, bgr_h = make_sequence_parallel(shard.sequence_of_rank(rank), F, max(1, min(16, ncpu // max(1, world))))
    disp_p = torch.from_numpy(disp_h).pin_memory()
    bgr_p = torch.from_numpy(bgr_h).pin_memory()
    cm = default_carmask((H, W), prefer_cwd=False)
    eng = RoadPath((H, W), step=args.step, trials=T, samples=S, carmask=cm, max_frames=F, device=dev)
    disp_d, bgr_d = disp_p.to(dev), bgr_p.to(dev)
    seed = 20260000 + rank

    def step():
        return eng.process(disp_d, bgr_d, prev=None, seed=seed, frame0=0)

    def barrier():
        shard.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    N.launches = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        for _ in range(args.steps):
            res = step()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = N.launches
    counts = res.counts.cpu().numpy().astype(np.float64)
    ref_counts = counts.copy()
    ref_mask = res.road_mask.cpu().numpy() if not args.no_e2e else None      # `res` is a view: later legs overwrite it
    planes_valid = float(res.planes[:, 7].sum().item())

    # ---- the one-stream variant of the step (RoadPath(overlap=False): whole-batch launches, nothing concurrent): measured
    # beside `value`, not part of it.  The per-kernel leg below runs this way too, so that a kernel's time is its own.
    eng.overlap = False
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(5):
        step()
    s1.record()
    torch.cuda.synchronize()
    one_stream_ms = s0.elapsed_time(s1) / 5
    # ---- the step without the 8-byte coloured cloud (RoadPath(emit_cloud=False): masks, planes, counts only -- what the
    # end-to-end entry point returns); beside `value`, not part of it
    eng_nc = RoadPath((H, W), step=args.step, trials=T, samples=S, carmask=cm, max_frames=F, device=dev, emit_cloud=False)
    for _ in range(3):
        rn = eng_nc.process(disp_d, bgr_d, prev=None, seed=seed, frame0=0)
    torch.cuda.synchronize()
    s0.record()
    for _ in range(5):
        rn = eng_nc.process(disp_d, bgr_d, prev=None, seed=seed, frame0=0)
    s1.record()
    torch.cuda.synchronize()
    nocloud_ms = s0.elapsed_time(s1) / 5
    nocloud_same = bool(torch.equal(rn.road_mask, res.road_mask) and torch.equal(rn.planes, res.planes) and torch.equal(rn.counts, res.counts))
    del eng_nc, rn
    torch.cuda.empty_cache()
    # ---- per-kernel device times (5 extra steps with event tracing on, after 2 untraced ones: the frees above leave the
    # GPU idle and the first launch after an idle gap is slow) -> roofline of the dominant kernel
    for _ in range(2):
        step()
    N.profile_enable(True)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    prof = N.profile_read()
    N.profile_enable(False)
    eng.overlap = True
    Nm, Mm, Km, Rm = counts.mean(axis=0)
    kern = {k: {"ms_per_launch": v[0] / v[1], "launches": v[1]} for k, v in prof.items()}
    tot = sum(v["ms_per_launch"] for v in kern.values())
    for k, v in kern.items():
        v["share"] = v["ms_per_launch"] / tot
        ab = algorithmic_bytes(k, eng.L, Nm, Mm, Km, Rm, T, S)
        if ab is not None:
            v["alg_bytes"] = ab * F
            v["gbs"] = ab * F / (v["ms_per_launch"] * 1e-3) / 1e9

    # ---- SURVEY 8(f)-1, the consumer of the road masks (sanitiseRoadImage): timed beside the path, not part of `value`
    next_rows = None
    if rank == 0 and not args.path_only:
        from stereo.vision_b200.pipeline import RoadCleaner
        cl = RoadCleaner((H, W), max_frames=F, device=dev)
        masks = res.road_mask
        for _ in range(3):
            cl.clean(masks)
            cl.obstacles()
        torch.cuda.synchronize()
        N.profile_enable(True)
        for _ in range(5):
            cl.clean(masks)
            cl.obstacles()
        torch.cuda.synchronize()
        cprof = N.profile_read()
        N.profile_enable(False)
        pts = float(cl.counts[:F].float().mean().item())
        calg = {"clean_band": 2.0 * H * W + 4.0 * ((W + 31) // 32) * H + 4.0 * H,     # mask in, cleaned bytes + bit rows + row counts out
                "mask_points": 4.0 * ((W + 31) // 32) * H + 4.0 * H + 8.0 * pts,       # bit rows + row counts in, (x,y) int32 pairs out
                "road_hull": 4.0 * ((W + 31) // 32) * H + 4.0 * H,                     # 8(f)-3: bit rows in, row spans out (latency-bound: one CTA per frame)
                "hull_masks": 4.0 * ((W + 31) // 32) * H + 4.0 * H + 2.0 * H * W}      # bit rows + spans in, hull mask + obstacle image out
        next_rows = {"row": "SURVEY 8(f)-1 sanitiseRoadImage and 8(f)-3 the obstacle stage on the step's %d road masks (not part of `value`)" % F,
                     "mean_points": pts, "status": int(cl.status.item()), "kernels": {}}
        for k, v in cprof.items():
            if k in calg:
                t_ms = v[0] / v[1]
                next_rows["kernels"][k] = {"ms_per_launch": round(t_ms, 6), "alg_bytes": calg[k] * F,
                                           "gbs": round(calg[k] * F / (t_ms * 1e-3) / 1e9, 3)}
        del cl
        # SURVEY 8(f)-2, the producer of the colour input (preProcessImages + greyscale) on the step's left images
        from stereo.vision_b200.pipeline import FrontEnd
        fe = FrontEnd((H, W), max_frames=F, device=dev)
        for _ in range(2):
            fe.process(bgr_d)
        torch.cuda.synchronize()
        N.profile_enable(True)
        for _ in range(3):
            fe.process(bgr_d)
        torch.cuda.synchronize()
        fprof = N.profile_read()
        N.profile_enable(False)
        falg = {"prep_gamma_gray": 7.0 * H * W + 1024.0,      # BGR in, corrected BGR + gray out, histogram
                "prep_apply_lut": 2.0 * H * W + 256.0}        # gray in, equalised gray out
        front = {"row": "SURVEY 8(f)-2 preProcessImages + greyscale on the step's %d left images (not part of `value`)" % F, "kernels": {}}
        for k, v in fprof.items():
            if k.startswith("prep_"):
                t_ms = v[0] / v[1]
                front["kernels"][k] = {"ms_per_launch": round(t_ms, 6)}
                if k in falg:
                    front["kernels"][k].update(alg_bytes=falg[k] * F, gbs=round(falg[k] * F / (t_ms * 1e-3) / 1e9, 3))
        next_rows = {"f1_road_mask_cleaning": next_rows, "f2_front_end": front}
        del fe
        # SURVEY 8(f)-4, the producer of the disparity input (functions.disparity = StereoSGBM + filterSpeckles + scaling)
        from stereo.vision_b200.pipeline import Matcher
        import stereo.vision_b200.synth as synth
        MF = 32                                            # all in flight: 9 GB of cost volumes
        pairs = [synth.make_stereo_pair(seed, t, (H, W)) for t in range(4)]
        gl = torch.from_numpy(np.stack([pairs[t % 4][0] for t in range(MF)])).to(dev)
        gr = torch.from_numpy(np.stack([pairs[t % 4][1] for t in range(MF)])).to(dev)
        mt = Matcher((H, W), max_frames=MF, frames_in_flight=MF, device=dev)
        mt.disparity(gl, gr)
        torch.cuda.synchronize()
        N.profile_enable(True)
        m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        m0.record()
        for _ in range(2):
            d8 = mt.disparity(gl, gr)
        m1.record()
        torch.cuda.synchronize()
        mprof = N.profile_read()
        N.profile_enable(False)
        m_ms = m0.elapsed_time(m1) / 2
        vol = 2.0 * H * (W - 128) * 128                   # one int16 cost volume of a frame
        malg = {"sgbm_prefilter": 2.0 * H * W * 17, "sgbm_hsum": 32.0 * H * W + vol, "sgbm_vsum": 2.25 * vol,
                "sgbm_path_v": 1.25 * vol, "sgbm_path_d1": 1.25 * vol, "sgbm_path_d3": 1.25 * vol, "sgbm_path_h": 1.25 * vol,
                "sgbm_final": 2 * vol + 2.0 * H * W, "sgbm_median": 4.0 * H * W, "speckle": 30.0 * H * W, "disparity_scale": 3.0 * H * W}
        match = {"row": "SURVEY 8(f)-4 functions.disparity (StereoSGBM 128 x block 21 + median + filterSpeckles + scaling) on %d synthetic "
                        "pairs, all in flight (not part of `value`)" % MF,
                 "frames_per_s": round(MF / (m_ms * 1e-3), 1), "ms_per_frame": round(m_ms / MF, 4),
                 "nonzero_fraction": round(float((d8 != 0).float().mean().item()), 4), "kernels": {}}
        for k, v in mprof.items():
            if k in malg:
                t_ms = v[0] / v[1]                         # per launch = all MF frames
                match["kernels"][k] = {"ms_per_launch": round(t_ms, 6), "alg_bytes": malg[k] * MF,
                                       "gbs": round(malg[k] * MF / (t_ms * 1e-3) / 1e9, 3)}
        if not args.no_cpu_baseline:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import sgbm_oracle as so                        # the checker, timed as the CPU port of this row (one frame, one core)
            t0 = time.perf_counter()
            want = so.disparity(pairs[0][0], pairs[0][1], 128, False)
            match["cpu_port"] = {"frames_per_s": round(1.0 / (time.perf_counter() - t0), 3), "cores": 1,
                                 "sample": "one 1024x544 pair through oracle/sgbm_oracle.c",
                                 "matches_gpu": bool(np.array_equal(want, d8[0].cpu().numpy()))}
        next_rows["f4_matcher"] = match
        del mt, gl, gr
        # every row together: camera images -> front end -> matcher -> path -> cleaning -> obstacles (stereovision.py:44-189)
        from stereo.vision_b200.pipeline import StereoPipeline
        torch.cuda.empty_cache()
        sp = StereoPipeline((H, W), max_frames=MF, trials=T, samples=S, step=args.step, carmask=cm, device=dev)
        four = [synth.make_stereo_bgr(seed, t, (H, W)) for t in range(4)]
        bl = torch.from_numpy(np.stack([four[t % 4][0] for t in range(MF)])).to(dev)
        br = torch.from_numpy(np.stack([four[t % 4][1] for t in range(MF)])).to(dev)
        sp.process(bl, br, seed=seed)
        torch.cuda.synchronize()
        m0.record()
        for _ in range(3):
            o = sp.process(bl, br, seed=seed)
        m1.record()
        torch.cuda.synchronize()
        w_ms = m0.elapsed_time(m1) / 3
        next_rows["whole_frame"] = {
            "row": "stereovision.py:44-189 for %d chained frames resident in HBM: front end on both views, matcher, path, cleaning, "
                   "obstacle stage (not part of `value`)" % MF,
            "frames_per_s": round(MF / (w_ms * 1e-3), 1), "ms_per_frame": round(w_ms / MF, 4), "kernels": {},
            "frames_with_plane": int((o["result"].planes[:MF, 7] > 0).sum().item()),
            "mean_road_points": float(o["counts"][:MF].float().mean().item())}
        del sp, bl, br

    # ---- result identity: this rank's hash of the step's results; rank 0 compares every rank's with the
    # single-GPU hashes of tests/golden/bench_sha.json (tools/make_bench_sha.py, one GPU, sequences 0-7)
    my_sha = shard.result_sha(res.road_mask, res.planes, res.counts) + ":" + shard.planes_sha(res.planes)[:16]

    # ---- the reference-identical mode: hypotheses = what `random.seed(ransac_seed(seq, t)); RANSAC(...)` draws in the
    # reference, replayed on the device (svb_ransac_replay_draws), everything else as in `value`
    parity = None
    if rank == 0 and not args.path_only:
        import stereo.vision_b200.synth as synth
        seeds = [synth.ransac_seed(shard.sequence_of_rank(rank), t) for t in range(F)]
        for _ in range(2):
            rp = eng.process(disp_d, bgr_d, prev=None, py_seeds=seeds)
        torch.cuda.synchronize()
        N.profile_enable(True)
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        for _ in range(3):
            rp = eng.process(disp_d, bgr_d, prev=None, py_seeds=seeds)
        p1.record()
        torch.cuda.synchronize()
        pprof = N.profile_read()
        N.profile_enable(False)
        p_ms = p0.elapsed_time(p1) / 3
        parity = {"row": "reference-identical hypotheses: CPython's random stream under random.seed(100003 seq + t) per frame, replayed on the "
                         "device; same %d-frame resident sequence" % F,
                  "frames_per_s": round(F / (p_ms * 1e-3), 1), "ms_per_step": round(p_ms, 4),
                  "kernels_ms": {k: round(v[0] / v[1], 5) for k, v in pprof.items()},
                  "result_sha": shard.result_sha(rp.road_mask, rp.planes, rp.counts),
                  "frames_with_plane": int((rp.planes[:, 7] > 0).sum().item())}
        eng._si = eng._ti = eng._replay_scratch = None          # 290 MB of index buffers back to the allocator
        del rp
        torch.cuda.empty_cache()

    # ---- the other configurations of BASELINE.json (single GPU; not part of `value`)
    configs = None
    if rank == 0 and not args.path_only:
        configs = extra_configs(args, torch, dev, N, cm, disp_d, bgr_d, seed)

    # ---- end to end: pinned host frames in, masks / planes / counts out, chunked + double buffered
    e2e = None
    if not args.no_e2e:
        with ClockSampler(local) as clk2:
            e2e = run_e2e(args, torch, dev, eng.camera, cm, disp_p, bgr_p, seed, ref_counts, ref_mask=ref_mask)
        e2e["clocks"] = clk2.summary()

    # ---- max over ranks
    ms_max, e2e_ms_max = shard.max_over_ranks([ms, e2e["ms"] if e2e else 0.0], dev)
    all_sha = shard.gather_strings(my_sha, dev)
    link_min = shard.min_over_ranks([e2e["link"]["h2d_gbs"], e2e["link"]["d2h_gbs"]], dev) if e2e else None
    out = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks \
            else (6650.0, "fallback (B200_PROFILING.md)")
        hbm_kernels = {k: v for k, v in kern.items() if "gbs" in v and k not in ("ransac_score", "road_clear")}
        dom = max(hbm_kernels, key=lambda k: hbm_kernels[k]["ms_per_launch"])
        top = max(kern, key=lambda k: kern[k]["ms_per_launch"])
        d = kern[dom]
        traffic, traffic_src = None, None                 # DRAM bytes per launch of that kernel from the committed ncu capture
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tj.get("frames") == F and dom in tj["kernels"]:
                traffic = tj["kernels"][dom]["dram_read_bytes"] + tj["kernels"][dom]["dram_write_bytes"]
                traffic_src = tj["source"]
        except Exception:
            pass
        out = {
            "metric": METRIC, "value": F * args.steps * world / (ms_max * 1e-3), "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8 in / f64 plane + reprojection arithmetic", "data": "synthetic",
            "config": {"workload": "config 2: %d-frame sequence per GPU, 1024x544, previous-disparity fill, step %d, "
                                   "%d RANSAC trials x %d samples, counter-based on-device draws" % (F, args.step, T, S),
                       "streams": "two half-batches on two streams (RoadPath's default for >= 192 frames; same results)",
                       "ms_per_step_one_stream": one_stream_ms,
                       "ms_per_step_without_cloud": nocloud_ms, "without_cloud_same_masks_planes_counts": nocloud_same,
                       "frames_per_step_per_gpu": F, "l2": "inputs (%.0f MB/sequence) exceed the 126 MB L2; no flush needed"
                                                          % ((disp_h.nbytes + bgr_h.nbytes) / 1e6),
                       "parallelism": "one independent sequence per GPU, no hot-path collective",
                       "mean_points": {"N": Nm, "M": Mm, "K": Km, "R": Rm}, "frames_with_plane": planes_valid},
            "gpu_launches": launches,
            "clocks": clk.summary(),
            "roofline": {"kernel": dom, "bound": "hbm", "achieved": d["gbs"], "peak": peak, "unit": "GB/s",
                         "frac": d["gbs"] / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "alg_bytes_per_launch": d["alg_bytes"], "ms_per_launch": d["ms_per_launch"],
                         "share_of_step": d["share"], "longest_kernel": top,
                         "timed": "whole-batch launches on one stream (5 steps with event tracing; `share_of_step` is of the one-stream step)"},
            "kernels": {k: {kk: (round(vv, 6) if isinstance(vv, float) else vv) for kk, vv in v.items()} for k, v in kern.items()},
        }
        gold, match = None, None
        try:
            gold = json.load(open(os.path.join(ROOT, "tests", "golden", "bench_sha.json")))
            same_cfg = (gold["frames"], gold["trials"], gold["samples"], gold["step"]) == (F, T, S, args.step)
            match = bool(same_cfg and all(all_sha[r] == gold["sha"][str(shard.sequence_of_rank(r))] for r in range(world)))
            if not same_cfg:
                match = None
        except Exception:
            pass
        out["result_sha"] = all_sha
        out["matches_single_gpu"] = match
        out["result_sha_note"] = ("per rank: sha256 of road masks | N,M,K,R | winning trials | plane flags, ':' first 16 hex of the sha256 of "
                                  "the planes' float64 bits; compared with tests/golden/bench_sha.json (one GPU, tools/make_bench_sha.py)")
        if parity:
            out["parity_mode"] = parity
        if configs:
            out["extra"] = {"configs": configs}
        if next_rows:
            for row in next_rows.values():
                for v in row["kernels"].values():
                    if "gbs" in v:
                        v["frac_of_peak"] = round(v["gbs"] / peak, 4)
            out["next_rows"] = next_rows
        if e2e:
            out["e2e"] = {"value": F * world / (e2e_ms_max * 1e-3), "unit": "frames/s",
                          "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"],
                          "ms_per_step": e2e_ms_max, "chunk_frames": args.chunk, "matches_resident_run": e2e["match"],
                          "api": "stereo.vision_b200.pipeline.SequenceStreamer(pack_mask=True).run (pinned host frames -> pinned host road masks as "
                                 "bit rows + planes + counts; the unpacked masks are checked against the resident run outside the timed region; "
                                 "uploaded: the lattice rows of the disparity and, once a chunk's planes are known, the lattice rows of the colour "
                                 "image that can hold a near-plane point -- nothing else of it is ever read)",
                          "upload": e2e["upload"], "with_every_lattice_row_of_the_colour_image": e2e["lattice"],
                          "gpu_launches_per_step": e2e["launches_per_step"], "clocks": e2e["clocks"],
                          "link": dict(e2e["link"], h2d_gbs_min_over_ranks=link_min[0], d2h_gbs_min_over_ranks=link_min[1]),
                          "h2d_gbs": e2e["h2d"] / (e2e_ms_max * 1e-3) / 1e9, "d2h_gbs": e2e["d2h"] / (e2e_ms_max * 1e-3) / 1e9}
            lk = out["e2e"]["link"]
            # the step moves h2d and d2h bytes concurrently over a full-duplex link: its floor is the slower direction alone
            floor_ms = max(e2e["h2d"] / (link_min[0] * 1e9), e2e["d2h"] / (link_min[1] * 1e9)) * 1e3
            out["e2e"]["link_floor_ms"] = floor_ms
            out["e2e"]["frac_of_link"] = floor_ms / e2e_ms_max
        if world == 1 and not args.no_cpu_baseline:
            procs = ncpu
            kind = _reference_kind(S, args.step)
            fpp = args.cpu_frames or (2 if kind == "reference" else 4)
            v, wall = cpu_throughput(kind, fpp, procs, T, S, args.step)
            out["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": procs, "kind": kind,
                                   "sample": cpu_sample_text(kind, procs, fpp, wall)}
            if kind == "reference":                    # the numpy port beside it, for the record (it is ~30x faster than the reference)
                v2, wall2 = cpu_throughput("port", 4, procs, T, S, args.step)
                out["cpu_baseline"]["port"] = {"value": v2, "sample": cpu_sample_text("port", procs, 4, wall2)}
        print(json.dumps(out))
    shard.finish()
    return out


def _timed(torch, fn, reps, warm=2):
    """mean device milliseconds of fn() over `reps` back-to-back calls"""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def extra_configs(args, torch, dev, N, cm, disp_d, bgr_d, seed):
    """BASELINE.json configs 1, 3 and 5 on one GPU (config 2 is `value`, config 4 the multi-GPU run)."""
    import random
    import stereo.vision_b200.synth as synth
    from stereo.vision_b200.pipeline import RoadPath
    T, S = args.trials, args.samples
    out = {}
    # ---- config 1: one frame, prev = None (single_frame.py:53): latency of one RoadPath.process call, host-synchronised
    one = RoadPath((H, W), step=2, trials=T, samples=S, carmask=cm, max_frames=1, device=dev)
    d1, c1 = disp_d[:1], bgr_d[:1]

    def latency(fn, reps=30):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e3)
        return float(np.median(ts))
    lat_dev = latency(lambda: one.process(d1, c1, seed=seed))
    lat_ref = latency(lambda: one.process(d1, c1, py_seeds=[synth.ransac_seed(0, 0)]))
    # the drop-in helper chain (stereovision.py:58-140 through stereo.vision_b200.functions, numpy in, numpy out)
    import stereo.vision_b200.functions as f
    f.carmask, f.ransac_samples = cm, S
    dn, cn = disp_d[0].cpu().numpy(), bgr_d[0].cpu().numpy()

    def helper_chain():
        random.seed(synth.ransac_seed(0, 0))
        disparity = f.disparityCleaning(dn, 'previous', None)
        masked = f.maskDisparity(disparity)
        points = f.projectDisparityTo3d(f.capDisparity(disparity), 128, cn)
        maskpoints = f.projectDisparityTo3d(masked, 128)
        normal, abc = f.RANSAC(maskpoints, T)
        diffs = f.calculatePointErrors(abc, points)
        points = f.computePlanarThreshold(points, diffs, 0.05)
        hist = f.calculateColourHistogram(points)
        points = f.filterPointsByHistogram(points, hist, 10)
        pp = np.array(f.project3DPointsTo2DImagePoints(points), np.int32).reshape((-1, 1, 2))
        return f.generatePointsAsImage(pp)

    def wall_ms(fn, reps):
        fn()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e3)
        return float(np.median(ts))
    f.ransac_replay = "device"
    chain_dev = wall_ms(helper_chain, 5)
    f.ransac_replay = "host"
    chain_host = wall_ms(helper_chain, 2)
    f.ransac_replay, f.ransac_samples = "device", 600
    out["config1_single_frame"] = {
        "row": "single_frame.py:53: one 1024x544 frame, prev = None, step 2, %d x %d; milliseconds per frame, host-synchronised" % (T, S),
        "fused_device_draws_ms": round(lat_dev, 4), "fused_reference_draws_ms": round(lat_ref, 4),
        "dropin_helper_chain_ms": round(chain_dev, 3), "dropin_helper_chain_host_replay_ms": round(chain_host, 3),
        "note": "fused = RoadPath.process(F=1) on resident inputs; drop-in = the twelve helper calls of stereovision.py:58-140 "
                "(numpy in, numpy out, Python's random stream replayed on the device / by the host loop)"}
    del one
    # ---- config 3: full density (step 1), colour attach, masked-cloud RANSAC
    F3 = min(60, disp_d.shape[0])
    e3 = RoadPath((H, W), step=1, trials=T, samples=S, carmask=cm, max_frames=F3, device=dev)
    ms3 = _timed(torch, lambda: e3.process(disp_d[:F3], bgr_d[:F3], seed=seed), 5)
    cnt3 = e3.counts[:F3].float().mean(dim=0).tolist()
    out["config3_step1"] = {"row": "step 1 (555 489 lattice sites per frame), %d chained frames resident, %d x %d" % (F3, T, S),
                            "frames_per_s": round(F3 / (ms3 * 1e-3), 1), "ms_per_frame": round(ms3 / F3, 5),
                            "mean_points": dict(zip("NMKR", [round(v, 1) for v in cnt3]))}
    del e3
    torch.cuda.empty_cache()
    # ---- config 5: 2048x1088 frames, RANSAC sweep
    shape5, F5 = (2 * H, 2 * W), 4
    from stereo.vision_b200.masks import default_carmask
    cm5 = default_carmask(shape5, prefer_cwd=False)
    frames5 = [synth.make_frame(900, t, shape5) for t in range(F5)]
    d5 = torch.from_numpy(np.stack([a for a, _ in frames5])).to(dev)
    c5 = torch.from_numpy(np.stack([b for _, b in frames5])).to(dev)
    sweep = []
    for (T5, S5) in ((600, 400), (2000, 1000), (10000, 4000)):
        e5 = RoadPath(shape5, step=2, trials=T5, samples=S5, carmask=cm5, max_frames=F5, device=dev)
        ms5 = _timed(torch, lambda: e5.process(d5, c5, seed=seed), 3, warm=1)
        N.profile_enable(True)
        e5.process(d5, c5, seed=seed)
        torch.cuda.synchronize()
        pr = N.profile_read()
        N.profile_enable(False)
        rs = pr["ransac_score"][0] / pr["ransac_score"][1]
        sweep.append({"trials": T5, "samples": S5, "ms_per_frame": round(ms5 / F5, 5), "ransac_score_ms_per_frame": round(rs / F5, 5),
                      "ransac_share": round(rs / sum(v[0] / v[1] for v in pr.values()), 3),
                      "sample_terms_per_s": round(F5 * T5 * S5 / (rs * 1e-3), 1),
                      "frames_with_plane": int((e5.planes[:F5, 7] > 0).sum().item())})
        del e5
        torch.cuda.empty_cache()
    out["config5_2048x1088"] = {"row": "2048x1088 frames (masks scaled x2), step 2, %d chained frames resident, device draws" % F5,
                                "sweep": sweep,
                                "note": "ransac_score is issue-bound (shared-memory gathers + 2 mad.wide + 3 fp64 per sample term): "
                                        "sample_terms_per_s is its throughput; profiles/ holds its warp-instruction count"}
    return out


def measure_link(torch, dev, h2d_bytes, d2h_bytes):
    """Plain pinned cudaMemcpyAsync bandwidth of this rank's link: one H2D copy of the step's upload size and one D2H
    copy of its download size, alone and both at once (all ranks run this at the same time: the ceiling e2e is held to)."""
    hb = torch.empty((h2d_bytes,), dtype=torch.uint8).pin_memory()
    db = torch.empty((h2d_bytes,), dtype=torch.uint8, device=dev)
    ho = torch.empty((d2h_bytes,), dtype=torch.uint8).pin_memory()
    do = torch.empty((d2h_bytes,), dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(up, down, reps=4):
        best = 1e30
        for _ in range(reps):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            s1.wait_event(a)
            s2.wait_event(a)
            if up:
                with torch.cuda.stream(s1):
                    db.copy_(hb, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    ho.copy_(do, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1)
            torch.cuda.current_stream().wait_stream(s2)
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        return best
    run(True, True, 1)
    up, down, both = run(True, False), run(False, True), run(True, True)
    return {"h2d_gbs": h2d_bytes / (up * 1e-3) / 1e9, "d2h_gbs": d2h_bytes / (down * 1e-3) / 1e9,
            "both_ms": both, "h2d_alone_ms": up, "d2h_alone_ms": down,
            "how": "one pinned cudaMemcpyAsync per direction of the step's byte counts, best of 4, CUDA events; every rank at the same time"}


def run_e2e(args, torch, dev, camera, cm, disp_p, bgr_p, seed, ref_counts, clk=None, ref_mask=None):
    """One sequence through the public host-to-host entry point (pipeline.SequenceStreamer.run): pinned
    HOST frames in, road masks + planes + counts in pinned host memory out; per chunk the H2D copy of
    the frames, the path and the D2H copy of the results run on three streams, double buffered."""
    from stereo.vision_b200.pipeline import SequenceStreamer
    F, C = args.frames, min(args.chunk, args.frames)
    sr = SequenceStreamer((H, W), step=args.step, trials=args.trials, samples=args.samples, carmask=cm, camera=camera,
                          chunk=C, device=dev, pack_mask=True)
    import stereo.vision_b200.shard as shard
    shard.barrier()                                       # every rank streams at the same time (rank 0 ran extra legs before)
    outs = [None, None]                                   # two result sets alternate: a step's results stay readable
    for i in range(2):                                    # while the next step's uploads have already begun
        outs[i] = sr.run(disp_p, bgr_p, seed=seed, out=outs[i])
    torch.cuda.synchronize()
    steps = max(3, min(args.steps, 10))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        outs[i % 2] = sr.run(disp_p, bgr_p, seed=seed, out=outs[i % 2])
    out = outs[(steps - 1) % 2]
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    match = bool(np.array_equal(out["counts"].numpy().astype(np.float64), ref_counts))
    if ref_mask is not None:                              # outside the timed region: the bit rows unpack to the resident run's masks
        match = match and bool(np.array_equal(sr.unpack_road_mask(out["road_bits"]), ref_mask))
    shard.barrier()
    link = measure_link(torch, dev, sr.h2d_bytes, sr.d2h_bytes)
    # beside it: the same with every lattice row of the colour image uploaded (upload="lattice", the r02s streamer)
    h2d, d2h, mode = sr.h2d_bytes, sr.d2h_bytes, sr.upload
    road_launches = sr.engines[0].road_launches
    del sr
    sl = SequenceStreamer((H, W), step=args.step, trials=args.trials, samples=args.samples, carmask=cm, camera=camera,
                          chunk=C, device=dev, pack_mask=True, upload="lattice")
    for i in range(2):
        outs[i] = sl.run(disp_p, bgr_p, seed=seed, out=outs[i])
    torch.cuda.synchronize()
    e0.record()
    for i in range(3):
        outs[i % 2] = sl.run(disp_p, bgr_p, seed=seed, out=outs[i % 2])
    e1.record()
    torch.cuda.synchronize()
    lattice = {"frames_per_s": F / (e0.elapsed_time(e1) / 3 * 1e-3), "h2d_bytes_per_step": sl.h2d_bytes,
               "same_results": bool(torch.equal(outs[0]["road_bits"], out["road_bits"]) and torch.equal(outs[0]["planes"], out["planes"]))}
    del sl
    return {"ms": ms, "h2d": h2d, "d2h": d2h, "match": match, "link": link, "upload": mode, "lattice": lattice,
            "launches_per_step": len(range(0, F, C)) * (road_launches + 4 + 1)}      # fill, write, score, finalize + svb_pack_mask


# ------------------------------------------------------------------ reference arm (CPU)
def run_reference(args):
    """The reference's CPU path for the same workload on the box's host cores: the reference's own helper chain
    (byte code under oracle/_ref/, built from /root/reference by oracle/build_ref.py; /root/reference itself does not
    exist on the GPU box), one independent sequence per host core; when that is not available, the oracle port.
    Each step is a bounded sample of the 300-frame workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    procs = os.cpu_count() or 1
    T, S = args.trials, args.samples
    kind = _reference_kind(S, args.step)
    fpp = args.cpu_frames or 2
    for _ in range(min(args.warmup, 1)):
        cpu_throughput(kind, 1, procs, T, S, args.step)
    vals, walls = [], []
    for _ in range(args.steps):
        v, wall = cpu_throughput(kind, fpp, procs, T, S, args.step)
        vals.append(v)
        walls.append(wall)
    v = float(np.mean(vals))
    sample = "each step = " + cpu_sample_text(kind, procs, fpp, float(np.mean(walls))) + " (of the %d-frame workload)" % args.frames
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * float(np.mean(walls)), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64 (python / numpy)", "data": "synthetic",
        "config": {"workload": "config 2: %d-frame sequence, 1024x544, previous-disparity fill, step %d, %d RANSAC "
                               "trials x %d samples (bounded sample per step)" % (args.frames, args.step, T, S)},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": procs, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

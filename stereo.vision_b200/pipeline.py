"""The fused, frame-batched path: stereovision.py:58-140 for F frames of one sequence per call.

`RoadPath.process` is the throughput entry point that sits beside the per-helper drop-in API
(functions.py of this package).  One call enqueues, on the current CUDA stream and without any host
synchronisation:

    svb_fill_sequence_counted    a1    temporal fill (the only cross-frame dependency, walked in-kernel)
                                       + the site counts of a4 (the bytes are in registers anyway)
    svb_project_compact_counted  a2-a4 write: cloud (8 B records) + masked cloud (4 B); offsets summed per block
    svb_ransac_score     a5-a7   T hypotheses x S samples per frame, one warp per hypothesis
    svb_ransac_finalize          first-minimum arg-min, winning plane, exact threshold cutoff
    svb_road_points      a8-a14  near-plane test + hue histogram, colour filter + reprojection + mask

All buffers are allocated once in the constructor (slabs of lattice-size records per frame).
"""
import ctypes as C

import numpy as np
import torch

from . import _native as N
from .camera import DEFAULT_CAMERA, Camera

KERNELS_PER_BATCH = 5      # fill+count, write, score, finalize, road cluster kernel


def _on_device(fn):
    """Run a method with the object's device current: the library launches on the current device's stream, so an
    engine built for cuda:1 must not launch on cuda:0 because the caller never called set_device."""
    import functools

    @functools.wraps(fn)
    def wrapped(self, *a, **kw):
        with torch.cuda.device(self.device):
            return fn(self, *a, **kw)
    return wrapped


def _as_cuda_u8(x, device):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    if not x.is_cuda:
        x = x.to(device, non_blocking=True)
    if x.dtype != torch.uint8:
        raise TypeError("expected uint8 data")
    return x.contiguous()


class BatchResult:
    """Device-resident outputs of one RoadPath.process call (views into the engine's buffers)."""

    def __init__(self, eng, F):
        self.F = F
        self.filled = eng.filled[:F]
        self.counts = eng.counts[:F]          # int32 [F,4]: N (cloud), M (masked cloud), K (near), R (road)
        self.planes = eng.planes[:F]          # float64 [F,8]: a,b,c,|abc|,cutoff,best_error,best_trial,valid
        self.errors = eng.errors[:F]          # float64 [F,T]
        self.hist = eng.hist[:F]              # uint32-as-int32 [F,1024]
        self.road_mask = eng.road_mask[:F]    # uint8 [F,H,W]
        self.cloud = eng.cloud[:F] if eng.emit_cloud else None     # int64 [F,L] (uint64 records)
        self.mcloud = eng.mcloud[:F]          # int32 [F,L] (uint32 records)
        self.overlay = None
        self.indices = None                   # (sample_idx, triple_idx) when the draws were explicit or replayed

    def abc(self):
        """Host float64 [F,3]; NaN rows where no plane was found (RANSAC -> (None, None))."""
        return self.planes[:, :3].cpu().numpy()

    def best_trial(self):
        return self.planes[:, 6].cpu().numpy().astype(np.int64)


class RoadPath:
    def __init__(self, shape=(544, 1024), step=2, trials=600, samples=600, point_threshold=0.05,
                 colour_threshold=10, carmask=None, camera: Camera = DEFAULT_CAMERA, max_frames=1,
                 device="cuda:0", use_reproj_table=True, road_path="sites", overlap=True, emit_cloud=True):
        if road_path not in ("sites", "records"):
            raise ValueError("road_path is 'sites' or 'records'")
        N.require_cuda()
        self.lib = N.load()
        self.device = torch.device(device)
        # overlap=True: a batch of >= 192 frames is processed as two half-batches on two streams, the hypothesis scoring of
        # one half beside the projection / road-point kernels of the other.  Same results, bit for bit (after the fill, a
        # frame depends on no other frame).  With the warp-per-hypothesis scorer (1024 threads, every register of the SM)
        # this gained nothing (0.585 against 0.574 ms per 300 frames); the thread-per-hypothesis scorer leaves a fifth of
        # the registers and its shared-memory gathers leave the DRAM idle: 0.466 against 0.493 ms.  Three ranges: 0.484,
        # four: 0.503, six: 0.58 - 0.68 (the launches get too small).
        self.overlap, self._side, self._scratch2 = bool(overlap), None, None
        # emit_cloud=False: the 8-byte coloured cloud (a4's first product) is neither written nor is the colour image
        # read by the projection -- for callers who want masks, planes and counts only (SequenceStreamer).  The masked
        # cloud (RANSAC's input) and N are produced as always; needs the road path that reads the disparity ("sites").
        self.emit_cloud = bool(emit_cloud)
        with torch.cuda.device(self.device):
            self._init(shape, step, trials, samples, point_threshold, colour_threshold, carmask, camera, max_frames,
                       use_reproj_table, road_path)

    def _init(self, shape, step, trials, samples, point_threshold, colour_threshold, carmask, camera, max_frames,
              use_reproj_table, road_path):
        self.H, self.W = int(shape[0]), int(shape[1])
        self.step, self.T, self.S = int(step), int(trials), int(samples)
        self.point_threshold = float(point_threshold)
        self.colour_threshold = int(colour_threshold)
        self.camera = camera
        self._cam = N.ccam(camera)
        self.LH = self.lib.svb_lattice_rows(self.H, self.step)
        self.LW = self.lib.svb_lattice_cols(self.W, self.step)
        self.L = self.LH * self.LW
        self.Fmax = int(max_frames)
        dev, F, H, W = self.device, self.Fmax, self.H, self.W
        if carmask is None:
            from .masks import default_carmask
            carmask = default_carmask((H, W))
        self.carmask = _as_cuda_u8(carmask, dev)
        if tuple(self.carmask.shape) != (H, W):
            raise ValueError("carmask shape %s != %s" % (tuple(self.carmask.shape), (H, W)))
        # most masked-cloud points a frame can have: lattice sites inside carmask (bounds the RANSAC staging)
        cm_host = self.carmask.cpu().numpy()
        self.max_m = int((cm_host[0:H - 1:self.step, 0:W - 1:self.step] != 0).sum())
        self.filled = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
        self.cloud = torch.empty((F if self.emit_cloud else 0, self.L), dtype=torch.int64, device=dev)
        self.mcloud = torch.empty((F, self.L), dtype=torch.int32, device=dev)
        self.counts = torch.zeros((F, 4), dtype=torch.int32, device=dev)
        self.errors = torch.empty((F, self.T), dtype=torch.float64, device=dev)
        self.planes = torch.zeros((F, N.PLANE_STRIDE), dtype=torch.float64, device=dev)
        self.hist = torch.zeros((F, N.HIST_STRIDE), dtype=torch.int32, device=dev)
        self.road_mask = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
        nb = max(self.lib.svb_project_scratch_bytes(F, H, W, self.step),
                 self.lib.svb_road_scratch_bytes(F, H, W, self.step),
                 self.lib.svb_road_sites_scratch_bytes(F, H, W, self.step))
        self.scratch = torch.empty(((nb + 15) // 16 * 16,), dtype=torch.uint8, device=dev)
        # reprojection displacement bits (built once; used only if every (coordinate, d) moves by 0 or -1)
        self.reproj_bits = torch.empty((self.lib.svb_reproject_bits_words(H, W),), dtype=torch.int32, device=dev)
        ok = torch.zeros((1,), dtype=torch.int32, device=dev)
        N.check(self.lib.svb_reproject_bits(H, W, C.byref(self._cam), N.ptr(self.reproj_bits), N.ptr(ok), N.stream_ptr()),
                "svb_reproject_bits")
        if int(ok.item()) != 1 or not use_reproj_table:
            self.reproj_bits = None
        # svb_road_points: one cluster kernel when a CTA's near list fits in shared memory, else the streamed path
        # (params, near, scatter kernels); without the displacement table the legacy two-pass kernels
        self.fused_road = self.reproj_bits is not None and ((self.L + 2047) // 2048 + 7) // 8 * 2048 * 4 <= 160 * 1024
        self.road_launches = 1 if self.fused_road else (3 if self.reproj_bits is not None else 2)
        # svb_road_points_sites: the same outputs from the filled disparity + colour image (1 byte per lattice site
        # instead of an 8-byte record); road_path = "records" keeps svb_road_points
        self.sites_road = (road_path == "sites" and self.reproj_bits is not None
                           and bool(self.lib.svb_road_sites_supported(H, W, self.step)))
        if self.sites_road:
            self.road_launches = 2
        if not self.emit_cloud and not self.sites_road:
            raise ValueError("emit_cloud=False needs road_path='sites' (the records path reads the cloud)")
        self._si = self._ti = self._replay_scratch = None      # reference-identical draws (py_seeds): allocated on first use

    # ------------------------------------------------------------------
    def _replay_buffers(self, F):
        if self._si is None:
            dev = self.device
            self._si = torch.empty((self.Fmax, self.T, self.S), dtype=torch.int32, device=dev)
            self._ti = torch.empty((self.Fmax, self.T, 3), dtype=torch.int32, device=dev)
            nb = int(self.lib.svb_ransac_replay_scratch_bytes(self.Fmax, self.S, max(self.max_m, 1)))
            self._replay_scratch = torch.empty((nb,), dtype=torch.uint8, device=dev)
            self._mt = torch.empty((self.Fmax, 625), dtype=torch.int32, device=dev)
        return self._si[:F], self._ti[:F]

    @_on_device
    def process(self, disp, bgr, prev=None, seed=0, frame0=0, indices=None, overlay=None,
                fill=True, py_seeds=None, py_stream=None, road=True) -> BatchResult:
        """Run the path on disp uint8 [F,H,W] + bgr uint8 [F,H,W,3].

        prev     : filled disparity of the frame before disp[0] (None at the start of a sequence);
        indices  : (sample_idx int32 [F,T,S], triple_idx int32 [F,T,3]) -> parity mode (explicit
                   draws, e.g. the replay of Python's `random`); None -> counter-based on-device
                   draws keyed by (seed, frame0 + t, trial);
        py_seeds : F integers -> REFERENCE-IDENTICAL draws: frame t picks the hypotheses the reference picks after
                   `random.seed(py_seeds[t])` (functions.py:286, 252-254), replayed on the device
                   (svb_ransac_replay_draws); the generators' end states are left in self.mt_state [F,625];
        py_stream: one integer seed, or a random.getstate() tuple -> ONE generator serves the F frames in order, as
                   after `random.seed(s)` before loop.py's frame loop (sequential on the device: one warp);
                   self.mt_state[0] then holds the end state (see mt_getstate) and can be passed on;
        overlay  : optional uint8 [F,H,W,3] CUDA tensor pre-filled with imgL; road sites get (0,255,0);
        fill     : False skips the temporal fill entirely (threshold_option other than 'previous');
        road     : False stops behind the plane (a1-a7); process_road(bgr) then runs a8-a14 on the same frames.  Needs
                   emit_cloud=False (then nothing before the road stage reads the colour image, and `bgr` may be None
                   here): SequenceStreamer(upload="needed") sends the colour rows only once the planes say which.
        """
        lib, dev = self.lib, self.device
        disp = _as_cuda_u8(disp, dev)
        if not road and (self.emit_cloud or not self.sites_road or overlay is not None):
            raise ValueError("road=False needs emit_cloud=False, the sites road path and no overlay")
        if bgr is None and road:
            raise ValueError("bgr is required unless road=False")
        bgr = _as_cuda_u8(bgr, dev) if bgr is not None else None
        if disp.dim() == 2:
            disp, bgr = disp[None], (bgr[None] if bgr is not None else None)
        F = disp.shape[0]
        if F > self.Fmax or tuple(disp.shape[1:]) != (self.H, self.W) or (bgr is not None and tuple(bgr.shape) != (F, self.H, self.W, 3)):
            raise ValueError("bad batch shape %s / %s" % (tuple(disp.shape), None if bgr is None else tuple(bgr.shape)))
        H, W = self.H, self.W
        res = BatchResult(self, F)
        # ---- draws: explicit index tensors, the reference's (replayed on the device per range below), or device draws
        si = ti = seeds_t = None
        chain = py_stream is not None
        replay = py_seeds is not None or chain
        if indices is not None:
            si = indices[0] if torch.is_tensor(indices[0]) else torch.from_numpy(np.ascontiguousarray(indices[0], dtype=np.int32))
            ti = indices[1] if torch.is_tensor(indices[1]) else torch.from_numpy(np.ascontiguousarray(indices[1], dtype=np.int32))
            si = si.to(dev).contiguous().view(F, self.T, self.S)
            ti = ti.to(dev).contiguous().view(F, self.T, 3)
        elif replay:
            si, ti = self._replay_buffers(F)
            if chain and not isinstance(py_stream, int):
                self.mt_setstate(py_stream)
            else:
                vals = [int(py_stream)] if chain else [int(v) for v in py_seeds]
                if len(vals) != (1 if chain else F) or min(vals) < 0 or max(vals) >= 1 << 64:
                    raise ValueError("py_seeds: one non-negative 64-bit integer per frame")
                seeds_t = torch.tensor([v - (1 << 64) if v >= 1 << 63 else v for v in vals], dtype=torch.int64).to(dev)
        res.indices = (si, ti) if si is not None else None
        ov = None
        if overlay is not None:
            ov = _as_cuda_u8(overlay, dev)
            res.overlay = ov
        # ---- frame ranges: one, or two half-batches on two side streams (self.overlap).  Only the fill chains frames
        # (functions.py:140-147), so it runs range after range on the caller's stream; everything behind it is per frame.
        cur = torch.cuda.current_stream(dev)
        parts = self.OVERLAP_PARTS if (self.overlap and F >= self.OVERLAP_PARTS * self.OVERLAP_MIN_FRAMES and not chain and road) else 1
        split = parts > 1
        ranges = [(F * k // parts, F * (k + 1) // parts) for k in range(parts)]
        if split and (self._side is None or len(self._side) < parts):
            self._side = tuple(torch.cuda.Stream(dev) for _ in range(parts))
            self._scratch2 = [torch.empty_like(self.scratch) for _ in range(parts - 1)]
        scratches = [self.scratch] + (self._scratch2[:parts - 1] if split else [])
        ready = []
        if fill:
            prev_t = _as_cuda_u8(prev, dev) if prev is not None else None
            if prev_t is not None and tuple(prev_t.shape) != (H, W):
                raise ValueError("prev is %s, the frames are %s" % (tuple(prev_t.shape), (H, W)))
            if prev_t is not None and prev_t.untyped_storage().data_ptr() == self.filled.untyped_storage().data_ptr():
                prev_t = prev_t.clone()     # result.filled[-1] of the last call: the fill kernel is about to overwrite it
            for (a, b), sc in zip(ranges, scratches):
                # a1 + the counting pass of a4 in one kernel; the scratch carries the counts to the projection
                N.check(lib.svb_fill_sequence_counted(N.ptr(disp[a:b]), N.ptr(prev_t if a == 0 else self.filled[a - 1]),
                                                      N.ptr(self.filled[a:b]), N.ptr(self.carmask), b - a, H, W, self.step,
                                                      N.ptr(sc), sc.numel(), C.c_void_p(cur.cuda_stream)), "svb_fill_sequence_counted")
                if split:
                    ready.append(torch.cuda.Event())
                    ready[-1].record(cur)
        else:
            self.filled[:F].copy_(disp)
            if split:
                ready = [torch.cuda.Event()] * parts
                ready[0].record(cur)
        for k, ((a, b), sc) in enumerate(zip(ranges, scratches)):
            stream = self._side[k] if split else cur
            if split:
                stream.wait_event(ready[k])
            st = C.c_void_p(stream.cuda_stream)
            self._stages_plane(a, b, bgr, si, ti, seeds_t, chain, replay, seed, frame0, fill, sc, st)
            if road:
                self._stages_road(a, b, bgr, ov, sc, st)
        if split:
            for k in range(parts):
                cur.wait_stream(self._side[k])
        N.count(len(ranges) * (KERNELS_PER_BATCH - 1 + (self.road_launches if road else 0) + (1 if replay else 0)) + (0 if fill else 2))
        self._road_pending = None if road else F
        return res

    @_on_device
    def process_road(self, bgr, overlay=None):
        """a8-a14 for the frames of the last process(..., road=False) call, on the current stream: bgr uint8 [F,H,W,3] (CUDA);
        only the lattice rows that hold near-plane points are read."""
        F = getattr(self, "_road_pending", None)
        if not F:
            raise RuntimeError("process_road() follows process(..., road=False)")
        if not (torch.is_tensor(bgr) and not bgr.is_cuda and bgr.is_pinned() and bgr.dtype == torch.uint8 and bgr.is_contiguous()):
            bgr = _as_cuda_u8(bgr, self.device)           # else: pinned host memory, read in place by the kernel (zero-copy)
        if tuple(bgr.shape) != (F, self.H, self.W, 3):
            raise ValueError("bgr is %s, expected %s" % (tuple(bgr.shape), (F, self.H, self.W, 3)))
        ov = _as_cuda_u8(overlay, self.device) if overlay is not None else None
        self._stages_road(0, F, bgr, ov, self.scratch, N.stream_ptr())
        N.count(self.road_launches)
        self._road_pending = None

    OVERLAP_MIN_FRAMES = 96                # per range: one stream 0.153 / 0.199 / 0.287 / 0.329 / 0.392 / 0.501 ms at 32 / 64 / 128 / 160 / 200 / 280
                                           # frames, two streams 0.261 / 0.264 / 0.295 / 0.327 / 0.377 / 0.476 ms
    OVERLAP_PARTS = 2

    def _stages_plane(self, a, b, bgr, si, ti, seeds_t, chain, replay, seed, frame0, fill, scratch, st):
        """a2-a7 of frames [a, b) of the batch, enqueued on stream `st`: clouds, draws, hypotheses, plane."""
        lib, H, W, n = self.lib, self.H, self.W, b - a
        cloud = self.cloud[a:b] if self.emit_cloud else None
        mcloud, counts, planes = self.mcloud[a:b], self.counts[a:b], self.planes[a:b]
        errors, src, bgr = self.errors[a:b], self.filled[a:b], (bgr[a:b] if bgr is not None else None)
        project = lib.svb_project_compact_counted if fill else lib.svb_project_compact
        N.check(project(N.ptr(src), N.ptr(self.carmask), N.ptr(bgr) if self.emit_cloud else None, n, H, W, self.step,
                        C.byref(self._cam), N.ptr(cloud), N.ptr(mcloud), N.ptr(counts), N.ptr(scratch), scratch.numel(), st),
                "svb_project_compact")
        if si is not None:
            si, ti = si[a:b], ti[a:b]
        if replay:
            per = (self._replay_scratch.numel() - 256) // self.Fmax // 4 * 4      # bytes of workspace per frame
            rs = self._replay_scratch[a * per:]                                   # ranges run concurrently: disjoint slices
            N.check(lib.svb_ransac_replay_draws(N.ptr(mcloud), 1, self.L, N.ptr(counts), 4, 1, n, self.T, self.S, max(self.max_m, 1),
                                                C.byref(self._cam), N.ptr(seeds_t[a:b] if (seeds_t is not None and not chain) else seeds_t),
                                                N.ptr(self._mt[a:b]), 1 if chain else 0, N.ptr(si), N.ptr(ti),
                                                N.ptr(rs), rs.numel(), st), "svb_ransac_replay_draws")
        N.check(lib.svb_ransac_score(N.ptr(mcloud), 1, self.L, N.ptr(counts), 4, 1, n, self.T, self.S, self.max_m, C.byref(self._cam),
                                     N.ptr(si), N.ptr(ti), seed, frame0 + a, 0, N.ptr(errors), st), "svb_ransac_score")
        N.check(lib.svb_ransac_finalize(N.ptr(mcloud), 1, self.L, N.ptr(counts), 4, 1, n, self.T, self.S, C.byref(self._cam),
                                        N.ptr(ti), seed, frame0 + a, N.ptr(errors), self.point_threshold, N.ptr(planes), st),
                "svb_ransac_finalize")

    def _stages_road(self, a, b, bgr, ov, scratch, st):
        """a8-a14 of frames [a, b): near-plane points, hue histogram and filter, reprojection, road mask (+ overlay)."""
        lib, H, W, n = self.lib, self.H, self.W, b - a
        cloud = self.cloud[a:b] if self.emit_cloud else None
        counts, planes, src, bgr = self.counts[a:b], self.planes[a:b], self.filled[a:b], bgr[a:b]
        ovr = ov[a:b] if ov is not None else None
        if self.sites_road:
            N.check(lib.svb_road_points_sites(N.ptr(src), N.ptr(bgr) if bgr.is_cuda else C.c_void_p(bgr.data_ptr()), N.ptr(planes), n, H, W, self.step, C.byref(self._cam),
                                              self.colour_threshold, N.ptr(self.reproj_bits), N.ptr(self.hist[a:b]),
                                              N.ptr(self.road_mask[a:b]), N.ptr(ovr), N.ptr(counts), N.ptr(scratch), scratch.numel(), st),
                    "svb_road_points_sites")
        else:
            N.check(lib.svb_road_points(N.ptr(cloud), N.ptr(counts), N.ptr(planes), n, H, W, self.step, C.byref(self._cam),
                                        self.colour_threshold, N.ptr(self.reproj_bits), N.ptr(self.hist[a:b]), N.ptr(self.road_mask[a:b]),
                                        N.ptr(ovr), N.ptr(counts), N.ptr(scratch), scratch.numel(), st), "svb_road_points")

    # the generator state of the reference-identical mode, in random.getstate() / setstate() form
    @property
    def mt_state(self):
        return self._mt

    def mt_setstate(self, state, row=0):
        self._replay_buffers(1)
        ver, words, _ = state
        if ver != 3 or len(words) != 625:
            raise ValueError("expected a random.getstate() tuple")
        w = np.array(words, dtype=np.uint32).view(np.int32)
        self._mt[row].copy_(torch.from_numpy(w))

    def mt_getstate(self, row=0):
        w = self._mt[row].cpu().numpy().view(np.uint32)
        return (3, tuple(int(v) for v in w), None)

    @_on_device
    def dump_indices(self, F, seed=0, frame0=0):
        """The counter-based draws the last process() call used (for the oracle): int32 [F,T,S], [F,T,3]."""
        si = torch.empty((F, self.T, self.S), dtype=torch.int32, device=self.device)
        ti = torch.empty((F, self.T, 3), dtype=torch.int32, device=self.device)
        N.check(self.lib.svb_ransac_dump_indices(N.ptr(self.mcloud), 1, self.L, N.ptr(self.counts), 4, 1, F, self.T,
                                                 self.S, C.byref(self._cam), seed, frame0, N.ptr(si), N.ptr(ti),
                                                 N.stream_ptr()), "svb_ransac_dump_indices")
        N.count(1)
        return si, ti


def near_row_ranges(planes, camera, H, W, step):
    """Per plane, the lattice rows of an H x W frame that CAN hold a near-plane point: (first, count) int32 arrays in
    lattice-row units, count 0 where no row can (or the plane is invalid).  planes: float64 [n, 8] rows of
    svb_ransac_finalize (a, b, c, |abc|, cutoff, ..., valid).

    With the plane in pixel form scaled by fB, s(x, y) = A x + B y + C is the disparity the plane itself has at (x, y)
    and P.abc - 1 = (s - d) / d (csrc/svb_sites.cu), so a point with disparity d in 1 .. 255 is near only if
    d (1 - cut) < s < d (1 + cut): a row whose largest s stays below min_d d (1 - cut), or whose smallest s stays above
    255 (1 + cut), holds none -- above the horizon of the plane the expected disparity is negative.  s is linear in y,
    so the rows that remain are one interval.  The margin (1e-3 of the magnitude of s plus 0.01) is a thousand times the
    float32 error band inside which the kernel lets the reference's float64 sequence decide."""
    pl = np.asarray(planes, dtype=np.float64).reshape(-1, 8)
    n = pl.shape[0]
    first, count = np.zeros(n, np.int32), np.zeros(n, np.int32)
    ok = (pl[:, 7] != 0) & np.isfinite(pl[:, :3]).all(axis=1) & np.isfinite(pl[:, 4])
    if not ok.any():
        return first, count
    q = np.where(ok[:, None], pl, 0.0)
    f, fB = camera.focal_px, camera.focal_px * camera.baseline_m
    A = fB * (q[:, 0] / f)
    B = fB * (q[:, 1] / f)
    Cc = fB * q[:, 2] - A * camera.centre_w - B * camera.centre_h
    cut = np.abs(q[:, 4])
    ys = np.arange(0, H - 1, step, dtype=np.float64)                        # lattice rows (functions.py:185)
    row = B[:, None] * ys[None, :] + Cc[:, None]
    smax = row + np.maximum(0.0, A * (W - 1))[:, None]
    smin = row + np.minimum(0.0, A * (W - 1))[:, None]
    m = 1e-3 * (np.abs(A) * W + np.abs(B) * H + np.abs(Cc)) + 0.01
    lo = np.minimum(1.0 - cut, 255.0 * (1.0 - cut)) - m
    hi = 255.0 * (1.0 + cut) + m
    can = (smax >= lo[:, None]) & (smin <= hi[:, None]) & ok[:, None]
    any_row = can.any(axis=1)
    lo_i = can.argmax(axis=1)
    hi_i = can.shape[1] - 1 - can[:, ::-1].argmax(axis=1)
    first[any_row] = lo_i[any_row]
    count[any_row] = (hi_i - lo_i + 1)[any_row]
    return first, count


class SequenceStreamer:
    """End-to-end entry point: a sequence that lives in HOST memory in, road masks / planes / counts in
    (pinned) host memory out -- what loop.py does frame by frame (loop.py:57-78), for a whole sequence.

    The sequence is cut into chunks of `chunk` frames.  Three streams overlap, per chunk, the host->device
    copy of the frames, the path (RoadPath.process, chained through the filled disparity of the previous
    chunk) and the device->host copy of the results; two engines / input buffers alternate.  Of the colour
    image only the lattice rows are sent (svb_upload_rows): the path never reads the others, and at step 2
    that is a quarter of all input bytes.  With upload="lattice" the same holds for the disparity:
    what run() returns -- road masks, planes, counts -- depends on the lattice rows of the (filled) disparity
    only (functions.py:185-186), so the rows in between are not sent either (they hold zeros on the device);
    upload="full" sends every disparity row, for a caller who also reads the engines' `filled` frames.
    upload="needed" goes one step further for the colour image, which is three quarters of the lattice bytes: with
    emit_cloud=False nothing reads it before the road stage, and the road stage reads it only at near-plane points
    (stereovision.py:97-103).  The disparity rows of a chunk go up first, the path runs up to the plane (a1-a7), the
    planes come back (1.9 KB), near_row_ranges() bounds the lattice rows that can hold a near point -- everything above
    the plane's horizon cannot -- and only that block of colour rows is sent (one 3-D copy, svb_upload_row_block)
    before the road stage (a8-a14) of the chunk runs, one chunk behind the next chunk's first half.  Same results;
    upload="pull" sends even less: the colour image is not copied at all, the road kernel reads the 16-pixel spans that
    hold near points straight from the caller's PINNED frames (zero-copy over the link); it is what upload="auto" (the
    default) picks whenever the engines allow it ("lattice" otherwise), falling back to "needed" for pageable frames.
    `h2d_bytes` / `d2h_bytes` report what one run() moves.
    """

    def __init__(self, shape=(544, 1024), step=2, trials=600, samples=600, carmask=None,
                 camera: Camera = DEFAULT_CAMERA, chunk=60, device="cuda:0", upload="auto", pack_mask=False, buffers=3, **kw):
        if upload not in ("auto", "lattice", "full", "needed", "pull"):
            raise ValueError("upload is 'auto', 'pull', 'needed', 'lattice' or 'full'")
        # buffers: engines / disparity buffers in rotation.  upload="pull" uses all of them (a chunk's disparity copy then never
        # waits for the road kernel of the chunk before last, which runs at the pace of the link: 85.8 k frames/s with three,
        # 81.4 k with two, 83.7 k with four); the other modes alternate the first two.
        # pack_mask=True: the road masks come back as bit rows ("road_bits" int32 [F, H * ceil(W / 32)], bit x & 31 of
        # word x >> 5 of a row = pixel x; unpack_road_mask() expands them) -- an eighth of the bytes over the link
        self.pack_mask = bool(pack_mask)
        kw.setdefault("emit_cloud", False)          # run() returns masks, planes and counts: nothing reads the cloud
        self.H, self.W = int(shape[0]), int(shape[1])
        self.step, self.chunk = int(step), int(chunk)
        self.device = torch.device(device)
        self.engines = [RoadPath(shape, step=step, trials=trials, samples=samples, carmask=carmask, camera=camera,
                                 max_frames=self.chunk, device=device, **kw) for _ in range(max(2, int(buffers)))]
        H, W, C_, dev = self.H, self.W, self.chunk, self.device
        nb = len(self.engines)
        self.din = [torch.zeros((C_, H, W), dtype=torch.uint8, device=dev) for _ in range(nb)]
        self.cin = [torch.zeros((C_, H, W, 3), dtype=torch.uint8, device=dev) for _ in range(2)]
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
        self.words = int(N.load().svb_bits_words(H, W))
        self.bits = [torch.empty((C_, self.words), dtype=torch.int32, device=dev) for _ in range(nb)] if self.pack_mask else None
        self.lib = N.load()
        self.d2h_bytes = 0
        self._last_cmp, self._last_out = [None] * nb, [None] * nb        # last use of buffer / engine k, across run() calls
        self._last_head = [None] * nb
        e = self.engines[0]
        can_need = not e.emit_cloud and e.sites_road and self.step >= 2 and H % self.step == 0
        if upload in ("needed", "pull") and not can_need:
            raise ValueError("upload='%s' needs emit_cloud=False, the sites road path and a step >= 2 that divides H" % upload)
        self.upload = ("pull" if can_need else "lattice") if upload == "auto" else upload        # "auto": the fewest bytes
        self._h2d_static, self._spans_host, self._spans_frames = 0, None, 0
        if self.upload in ("needed", "pull"):
            self._pl_host = [torch.empty((C_, N.PLANE_STRIDE), dtype=torch.float64).pin_memory() for _ in range(2)]

    def _upload_rows(self, dst, src, n, row):
        """Lattice rows (`row` bytes each) of src [n,H,...] (host) -> the same rows of dst (device), on the current stream."""
        H, st = self.H, self.step
        LH = self.engines[0].LH
        if st == 1 or not src.is_contiguous():
            dst[:n].copy_(src, non_blocking=True)
            return n * H * row
        sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if H % st == 0:                                   # rows y = 0, st, 2 st, ... run on across frame boundaries
            rows = n * (H // st)
            N.check(self.lib.svb_upload_rows(C.c_void_p(src.data_ptr()), N.ptr(dst), st * row, row, rows, sp), "svb_upload_rows")
            return rows * row
        for f in range(n):
            N.check(self.lib.svb_upload_rows(C.c_void_p(src[f].data_ptr()), C.c_void_p(dst[f].data_ptr()), st * row, row, LH, sp),
                    "svb_upload_rows")
        return n * LH * row

    @_on_device
    def run(self, disp, bgr, seed=0, out=None, py_seeds=None):
        """disp uint8 [F,H,W], bgr uint8 [F,H,W,3]: CPU tensors (pin them for asynchronous copies).
        py_seeds: F integers -> the reference's own draws per frame (RoadPath.process), else device draws keyed by `seed`.
        Returns dict(road_mask uint8 [F,H,W], planes float64 [F,8], counts int32 [F,4]) of pinned CPU tensors
        (pass the previous result as `out` to reuse them).  Asynchronous: the current stream waits for the
        result; synchronise (or use run_sync) before reading it on the host -- and leave disp / bgr alone until then:
        the copies are asynchronous, and with upload="pull" the road kernel reads the pinned colour frames in place."""
        if isinstance(disp, np.ndarray):
            disp = torch.from_numpy(disp)
        if isinstance(bgr, np.ndarray):
            bgr = torch.from_numpy(bgr)
        F, H, W, Cn = disp.shape[0], self.H, self.W, self.chunk
        if tuple(disp.shape) != (F, H, W) or tuple(bgr.shape) != (F, H, W, 3):
            raise ValueError("bad sequence shape %s / %s" % (tuple(disp.shape), tuple(bgr.shape)))
        mkey = "road_bits" if self.pack_mask else "road_mask"
        if out is None or mkey not in out or out[mkey].shape[0] != F:
            out = {"planes": torch.empty((F, N.PLANE_STRIDE), dtype=torch.float64).pin_memory(),
                   "counts": torch.empty((F, 4), dtype=torch.int32).pin_memory()}
            if self.pack_mask:
                out["road_bits"] = torch.empty((F, self.words), dtype=torch.int32).pin_memory()
            else:
                out["road_mask"] = torch.empty((F, H, W), dtype=torch.uint8).pin_memory()
        s_in, s_cmp, s_out = self.s_in, self.s_cmp, self.s_out
        chunks = [(a, min(F, a + Cn)) for a in range(0, F, Cn)]
        if self.upload == "pull" and disp.is_contiguous() and bgr.is_contiguous() and bgr.is_pinned():
            return self._run_pull(disp, bgr, seed, out, py_seeds, chunks)
        if self.upload in ("needed", "pull") and disp.is_contiguous() and bgr.is_contiguous():
            return self._run_needed(disp, bgr, seed, out, py_seeds, chunks)          # pageable colour frames cannot be read in place
        ev_in, ev_cmp, ev_out = {}, {}, {}
        h2d = 0
        # The inputs live in host memory, so the copy stream does not wait for the caller's stream: the first
        # uploads of this call overlap the tail of the previous one.  Buffers and engines are guarded by the
        # events of their last use, within and across calls.
        for i, (a, b) in enumerate(chunks):
            k, n = i % 2, b - a
            busy_cmp = ev_cmp[i - 2] if i >= 2 else self._last_cmp[k]
            busy_out = ev_out[i - 2] if i >= 2 else self._last_out[k]
            with torch.cuda.stream(s_in):
                if busy_cmp is not None:
                    s_in.wait_event(busy_cmp)                 # input buffer k is free once its last chunk was computed
                if self.upload == "lattice":
                    h2d += self._upload_rows(self.din[k], disp[a:b], n, W)
                else:
                    self.din[k][:n].copy_(disp[a:b], non_blocking=True)
                    h2d += n * H * W
                h2d += self._upload_rows(self.cin[k], bgr[a:b], n, W * 3)
                ev_in[i] = torch.cuda.Event()
                ev_in[i].record(s_in)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(ev_in[i])
                if busy_out is not None:
                    s_cmp.wait_event(busy_out)                # engine k's outputs were read back
                prev = self.engines[1 - k].filled[chunks[i - 1][1] - chunks[i - 1][0] - 1] if i else None
                r = self.engines[k].process(self.din[k][:n], self.cin[k][:n], prev=prev, seed=seed, frame0=a,
                                            py_seeds=None if py_seeds is None else py_seeds[a:b])
                if self.pack_mask:
                    N.check(self.lib.svb_pack_mask(N.ptr(r.road_mask), n, H, W, N.ptr(self.bits[k]), None,
                                                   C.c_void_p(s_cmp.cuda_stream)), "svb_pack_mask")
                    N.count(1)
                ev_cmp[i] = torch.cuda.Event()
                ev_cmp[i].record(s_cmp)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_cmp[i])
                if self.pack_mask:
                    out["road_bits"][a:b].copy_(self.bits[k][:n], non_blocking=True)
                else:
                    out["road_mask"][a:b].copy_(r.road_mask, non_blocking=True)
                out["planes"][a:b].copy_(r.planes, non_blocking=True)
                out["counts"][a:b].copy_(r.counts, non_blocking=True)
                ev_out[i] = torch.cuda.Event()
                ev_out[i].record(s_out)
            self._last_cmp[k], self._last_out[k] = ev_cmp[i], ev_out[i]
        torch.cuda.current_stream(self.device).wait_stream(s_out)   # the caller's stream sees the results
        self._h2d_static, self._spans_frames = int(h2d), 0
        self.d2h_bytes = int(sum(t.numel() * t.element_size() for t in out.values()))
        self._last = out
        return out

    def _run_needed(self, disp, bgr, seed, out, py_seeds, chunks):
        """run() with upload="needed": per chunk, head = disparity rows up + a1-a7 + planes down, tail = the colour rows the
        planes allow + a8-a14 + results down; the tail of chunk i is issued behind the head of chunk i + 1."""
        H, W, st = self.H, self.W, self.step
        s_in, s_cmp, s_out = self.s_in, self.s_cmp, self.s_out
        ev_head, ev_cmp, ev_out, res = {}, {}, {}, {}
        h2d = 0
        cam = self.engines[0].camera

        def head(i):
            (a, b), k = chunks[i], i % 2
            n = b - a
            busy_head = ev_head[i - 2] if i >= 2 else self._last_head[k]
            busy_out = ev_out[i - 2] if i >= 2 else self._last_out[k]
            with torch.cuda.stream(s_in):
                if busy_head is not None:
                    s_in.wait_event(busy_head)                # disparity buffer k is free once its last chunk was filled
                up = self._upload_rows(self.din[k], disp[a:b], n, W)
                ev_d = torch.cuda.Event()
                ev_d.record(s_in)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(ev_d)
                if busy_out is not None:
                    s_cmp.wait_event(busy_out)                # engine k's outputs were read back
                prev = self.engines[1 - k].filled[chunks[i - 1][1] - chunks[i - 1][0] - 1] if i else None
                res[i] = self.engines[k].process(self.din[k][:n], None, prev=prev, seed=seed, frame0=a,
                                                 py_seeds=None if py_seeds is None else py_seeds[a:b], road=False)
                self._pl_host[k][:n].copy_(res[i].planes, non_blocking=True)
                ev_head[i] = torch.cuda.Event()
                ev_head[i].record(s_cmp)
            self._last_head[k] = ev_head[i]
            return up

        def tail(i):
            (a, b), k = chunks[i], i % 2
            n = b - a
            ev_head[i].synchronize()                          # the chunk's planes are in host memory
            # one block for the chunk: the union of the frames' ranges.  A 2-D copy per frame with its own range moves 7 %
            # fewer bytes and is slower (53 k against 65 k frames/s: thirty small copies per chunk keep the link at 40 GB/s);
            # a 3-D copy per group of 10 frames moves 3 % fewer bytes at 3 % less link rate (66.5 k against 66.3 k frames/s)
            f0, cnt = near_row_ranges(self._pl_host[k][:n].numpy(), cam, H, W, st)
            has = cnt > 0
            first = int(f0[has].min()) if has.any() else 0
            rows = int((f0[has] + cnt[has]).max()) - first if has.any() else 0
            busy_cmp = ev_cmp[i - 2] if i >= 2 else self._last_cmp[k]
            with torch.cuda.stream(s_in):
                if busy_cmp is not None:
                    s_in.wait_event(busy_cmp)                 # colour buffer k is free once its last chunk's road stage ran
                if rows:
                    N.check(self.lib.svb_upload_row_block(C.c_void_p(bgr[a:b].data_ptr()), N.ptr(self.cin[k]), st * W * 3, W * 3, H // st,
                                                          first, rows, n, C.c_void_p(s_in.cuda_stream)), "svb_upload_row_block")
                ev_c = torch.cuda.Event()
                ev_c.record(s_in)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(ev_c)
                self.engines[k].process_road(self.cin[k][:n])
                r = res.pop(i)
                if self.pack_mask:
                    N.check(self.lib.svb_pack_mask(N.ptr(r.road_mask), n, H, W, N.ptr(self.bits[k]), None,
                                                   C.c_void_p(s_cmp.cuda_stream)), "svb_pack_mask")
                    N.count(1)
                ev_cmp[i] = torch.cuda.Event()
                ev_cmp[i].record(s_cmp)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_cmp[i])
                if self.pack_mask:
                    out["road_bits"][a:b].copy_(self.bits[k][:n], non_blocking=True)
                else:
                    out["road_mask"][a:b].copy_(r.road_mask, non_blocking=True)
                out["planes"][a:b].copy_(r.planes, non_blocking=True)
                out["counts"][a:b].copy_(r.counts, non_blocking=True)
                ev_out[i] = torch.cuda.Event()
                ev_out[i].record(s_out)
            self._last_cmp[k], self._last_out[k] = ev_cmp[i], ev_out[i]
            return n * rows * W * 3

        for i in range(len(chunks)):
            h2d += head(i)
            if i:
                h2d += tail(i - 1)
        h2d += tail(len(chunks) - 1)
        torch.cuda.current_stream(self.device).wait_stream(s_out)   # the caller's stream sees the results
        self._h2d_static, self._spans_frames = int(h2d), 0
        self.d2h_bytes = int(sum(t.numel() * t.element_size() for t in out.values())) + disp.shape[0] * N.PLANE_STRIDE * 8
        self._last = out
        return out

    @property
    def h2d_bytes(self):
        """Bytes the last run() moved host -> device (upload="pull": read after synchronising, it includes what the road
        kernel read in place: 48 bytes per colour span it reports in slot SVB_HIST_SPANS of the frames' histogram rows)."""
        pulled = 48 * int(self._spans_host[:self._spans_frames].sum()) if self._spans_frames else 0
        return self._h2d_static + pulled

    def _run_pull(self, disp, bgr, seed, out, py_seeds, chunks):
        """run() with upload="pull": the disparity's lattice rows go up by DMA; the colour image is not copied at all -- the
        road kernel reads the 16-pixel spans that hold near-plane points straight from the caller's pinned frames (mapped
        into the device's address space like every pinned allocation), 41 % of the lattice's colour bytes on the bench
        sequence where the rows below the horizon are 54 %.  One pass per chunk, no wait on the host."""
        H, W = self.H, self.W
        F = disp.shape[0]
        s_in, s_cmp, s_out = self.s_in, self.s_cmp, self.s_out
        if self._spans_host is None or self._spans_host.shape[0] < F:
            self._spans_host = torch.zeros((F,), dtype=torch.int32).pin_memory()
        ev_cmp, ev_out = {}, {}
        h2d, nb = 0, len(self.engines)
        for i, (a, b) in enumerate(chunks):
            k, n = i % nb, b - a
            busy_cmp = ev_cmp[i - nb] if i >= nb else self._last_cmp[k]
            busy_out = ev_out[i - nb] if i >= nb else self._last_out[k]
            with torch.cuda.stream(s_in):
                if busy_cmp is not None:
                    s_in.wait_event(busy_cmp)                 # disparity buffer k is free once its last chunk was computed
                h2d += self._upload_rows(self.din[k], disp[a:b], n, W)
                ev_in = torch.cuda.Event()
                ev_in.record(s_in)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(ev_in)
                if busy_out is not None:
                    s_cmp.wait_event(busy_out)                # engine k's outputs were read back
                prev = self.engines[(i - 1) % nb].filled[chunks[i - 1][1] - chunks[i - 1][0] - 1] if i else None
                r = self.engines[k].process(self.din[k][:n], None, prev=prev, seed=seed, frame0=a,
                                            py_seeds=None if py_seeds is None else py_seeds[a:b], road=False)
                self.engines[k].process_road(bgr[a:b])        # pinned host frames, read in place
                if self.pack_mask:
                    N.check(self.lib.svb_pack_mask(N.ptr(r.road_mask), n, H, W, N.ptr(self.bits[k]), None,
                                                   C.c_void_p(s_cmp.cuda_stream)), "svb_pack_mask")
                    N.count(1)
                ev_cmp[i] = torch.cuda.Event()
                ev_cmp[i].record(s_cmp)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_cmp[i])
                if self.pack_mask:
                    out["road_bits"][a:b].copy_(self.bits[k][:n], non_blocking=True)
                else:
                    out["road_mask"][a:b].copy_(r.road_mask, non_blocking=True)
                out["planes"][a:b].copy_(r.planes, non_blocking=True)
                out["counts"][a:b].copy_(r.counts, non_blocking=True)
                self._spans_host[a:b].copy_(r.hist[:, N.HIST_SPANS], non_blocking=True)
                ev_out[i] = torch.cuda.Event()
                ev_out[i].record(s_out)
            self._last_cmp[k], self._last_out[k] = ev_cmp[i], ev_out[i]
        torch.cuda.current_stream(self.device).wait_stream(s_out)   # the caller's stream sees the results
        self._h2d_static, self._spans_frames = int(h2d), F
        self.d2h_bytes = int(sum(t.numel() * t.element_size() for t in out.values())) + 4 * F
        self._last = out
        return out

    def unpack_road_mask(self, road_bits):
        """uint8 [F,H,W] (0 / 255) of the bit rows run() returns with pack_mask=True (host side, numpy)"""
        b = road_bits.numpy() if torch.is_tensor(road_bits) else np.asarray(road_bits)
        F = b.shape[0]
        wpr = (self.W + 31) // 32
        bits = np.unpackbits(b.reshape(F, self.H, wpr).view(np.uint8), axis=2, bitorder="little")[:, :, :self.W]
        return bits * np.uint8(255)

    def run_sync(self, disp, bgr, seed=0, out=None, py_seeds=None):
        out = self.run(disp, bgr, seed=seed, out=out, py_seeds=py_seeds)
        torch.cuda.synchronize(self.device)
        return out


class RoadCleaner:
    """The consumer of the road masks: sanitiseRoadImage (functions.py:346-366, called at
    stereovision.py:156) for F frames per call -- closing, two erosions, road_threshold_mask, closing
    on bit-packed rows, then the row-major (x,y) list of what is left.

    clean(masks) enqueues two kernels on the current stream (svb_clean_road, svb_mask_points) and
    returns device tensors: cleaned [F,H,W] uint8, counts [F] int32, points [F,cap,2] int32 of which
    the first counts[f] rows of frame f are valid.  Buffers are allocated once.
    """

    def __init__(self, shape=(544, 1024), threshold_mask=None, max_frames=300, cap=None, device="cuda:0",
                 want_bytes=True):
        from .masks import default_road_threshold_mask
        N.require_cuda()
        self.H, self.W = int(shape[0]), int(shape[1])
        self.device = torch.device(device)
        self.lib = N.load()
        self.F = int(max_frames)
        H, W, F, dev = self.H, self.W, self.F, self.device
        self.words = int(self.lib.svb_bits_words(H, W))
        tm = default_road_threshold_mask((H, W)) if threshold_mask is None else np.ascontiguousarray(threshold_mask, dtype=np.uint8)
        if tm.shape != (H, W):
            raise ValueError("threshold mask is %s, frames are %s" % (tm.shape, (H, W)))
        self.cap = int(cap) if cap is not None else int((tm != 0).sum())        # nothing survives outside the mask
        self.tbits = torch.empty((self.words,), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            t = torch.from_numpy(tm).to(dev)
            N.check(self.lib.svb_pack_mask(N.ptr(t), 1, H, W, N.ptr(self.tbits), 0, N.stream_ptr()), "svb_pack_mask")
            torch.cuda.current_stream().synchronize()
        self.cleaned = torch.empty((F, H, W), dtype=torch.uint8, device=dev) if want_bytes else None
        self.bits = torch.empty((F, self.words), dtype=torch.int32, device=dev)
        self.rowcnt = torch.empty((F, H), dtype=torch.int32, device=dev)
        self.counts = torch.empty((F,), dtype=torch.int32, device=dev)
        self.points = torch.empty((F, max(self.cap, 1), 2), dtype=torch.int32, device=dev)
        self.status = torch.zeros((1,), dtype=torch.int32, device=dev)

    @_on_device
    def clean(self, masks, points=True):
        masks = _as_cuda_u8(masks, self.device)
        if masks.dim() == 2:
            masks = masks[None]
        F = masks.shape[0]
        if tuple(masks.shape[1:]) != (self.H, self.W) or F > self.F:
            raise ValueError("expected up to %d masks of %dx%d, got %s" % (self.F, self.H, self.W, tuple(masks.shape)))
        st = N.stream_ptr()
        N.check(self.lib.svb_clean_road(N.ptr(masks), N.ptr(self.tbits), F, self.H, self.W,
                                        N.ptr(self.cleaned) if self.cleaned is not None else 0, N.ptr(self.bits),
                                        N.ptr(self.rowcnt), N.ptr(self.status), st), "svb_clean_road")
        N.count(1)
        if points:
            N.check(self.lib.svb_mask_points(N.ptr(self.bits), N.ptr(self.rowcnt), F, self.H, self.W, self.cap,
                                             N.ptr(self.points), N.ptr(self.counts), st), "svb_mask_points")
            N.count(1)
        cleaned = self.cleaned[:F] if self.cleaned is not None else None
        self._last = F
        return cleaned, self.counts[:F], self.points[:F]

    @_on_device
    def obstacles(self, masks=True, hull_cap=512):
        """The obstacle stage (stereovision.py:170-200) of the frames of the last clean(): svb_road_hull,
        two kernels (one without masks).  Returns a dict of device tensors: hull [F,hull_cap,2] int32 (x,y) with hull_n [F]
        valid rows each (cv2.convexHull's vertices and orientation), centre [F,2] int32 (getCenterPoint;
        -1,-1 for an empty image) and, with masks=True, hull_mask and obstacles uint8 [F,H,W]."""
        F = getattr(self, "_last", 0)
        if F < 1:
            raise RuntimeError("obstacles() follows clean()")
        dev = self.device
        if getattr(self, "_hull_cap", None) != int(hull_cap):
            self._hull_cap = int(hull_cap)
            self.hull = torch.empty((self.F, self._hull_cap, 2), dtype=torch.int32, device=dev)
            self.hull_n = torch.empty((self.F,), dtype=torch.int32, device=dev)
            self.centre = torch.empty((self.F, 2), dtype=torch.int32, device=dev)
        if masks and getattr(self, "hull_mask", None) is None:
            self.hull_mask = torch.empty((self.F, self.H, self.W), dtype=torch.uint8, device=dev)
            self.obstacle_mask = torch.empty((self.F, self.H, self.W), dtype=torch.uint8, device=dev)
        sb = int(self.lib.svb_road_hull_scratch_bytes(self.F, self.H))
        if masks and getattr(self, "_hull_scratch", None) is None:
            self._hull_scratch = torch.empty((sb,), dtype=torch.uint8, device=dev)
        N.check(self.lib.svb_road_hull(N.ptr(self.bits), F, self.H, self.W, self._hull_cap,
                                       N.ptr(self.hull), N.ptr(self.hull_n), N.ptr(self.centre),
                                       N.ptr(self.hull_mask) if masks else None, N.ptr(self.obstacle_mask) if masks else None,
                                       N.ptr(self._hull_scratch) if masks else None, sb if masks else 0,
                                       N.stream_ptr()), "svb_road_hull")
        N.count(2 if masks else 1)
        out = {"hull": self.hull[:F], "hull_n": self.hull_n[:F], "centre": self.centre[:F]}
        if masks:
            out["hull_mask"], out["obstacles"] = self.hull_mask[:F], self.obstacle_mask[:F]
        return out


class FrontEnd:
    """The producer of the path's colour input: preProcessImages + greyscale (functions.py:80-96, called at
    stereovision.py:44,46) for F images per call -- gamma table on every BGR byte, gray of the corrected
    colours, per-image histogram, equalizeHist's table, equalised gray.  Three kernels per call
    (svb_gamma_gray_hist, svb_equalize_lut, svb_apply_lut_frames); buffers are allocated once."""

    def __init__(self, shape=(544, 1024), gamma=1.4, max_frames=300, device="cuda:0"):
        N.require_cuda()
        self.H, self.W, self.F = int(shape[0]), int(shape[1]), int(max_frames)
        self.device = torch.device(device)
        self.lib = N.load()
        inv = 1.0 / gamma
        self.table = np.ascontiguousarray(np.array([((i / 255.0) ** inv) * 255 for i in np.arange(0, 256)]).astype("uint8"))
        self._tp = self.table.ctypes.data_as(C.c_void_p)
        F, H, W, dev = self.F, self.H, self.W, self.device
        self.bgr = torch.empty((F, H, W, 3), dtype=torch.uint8, device=dev)
        self.gray = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
        self.eq = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
        self.hist = torch.empty((F, 256), dtype=torch.int32, device=dev)
        self.lut = torch.empty((F, 256), dtype=torch.uint8, device=dev)

    @_on_device
    def process(self, bgr):
        """bgr uint8 [F,H,W,3] -> (gamma-corrected bgr [F,H,W,3], equalised gray [F,H,W]) device tensors."""
        bgr = _as_cuda_u8(bgr, self.device)
        if bgr.dim() == 3:
            bgr = bgr[None]
        F = bgr.shape[0]
        if tuple(bgr.shape[1:]) != (self.H, self.W, 3) or F > self.F:
            raise ValueError("expected up to %d images of %dx%dx3, got %s" % (self.F, self.H, self.W, tuple(bgr.shape)))
        st = N.stream_ptr()
        N.check(self.lib.svb_gamma_gray_hist(N.ptr(bgr), F, self.H, self.W, self._tp, N.ptr(self.bgr), N.ptr(self.gray),
                                             N.ptr(self.hist), st), "svb_gamma_gray_hist")
        N.check(self.lib.svb_equalize_lut(N.ptr(self.hist), F, N.ptr(self.lut), st), "svb_equalize_lut")
        N.check(self.lib.svb_apply_lut_frames(N.ptr(self.gray), F, self.H * self.W, N.ptr(self.lut), N.ptr(self.eq), st),
                "svb_apply_lut_frames")
        N.count(3)
        return self.bgr[:F], self.eq[:F]


class Matcher:
    """The producer of the path's disparity input: functions.disparity (functions.py:104-128, called at
    stereovision.py:56) for F rectified gray pairs per call -- the module's StereoSGBM_create(0, 128, 21)
    (functions.py:26) with its 3x3 median, filterSpeckles(.., 0, 4000, max_disparity - 5), threshold, /16 and
    the x (256 / max_disparity) scaling that makes the uint8 disparity the path consumes.  The cost volumes
    (250 MB per 1024x544 frame) live in a scratch buffer sized for `frames_in_flight` frames; a longer batch is
    walked in chunks inside svb_sgbm_compute."""

    def __init__(self, shape=(544, 1024), max_disparity=128, block=21, max_frames=32, frames_in_flight=8, device="cuda:0"):
        N.require_cuda()
        self.H, self.W, self.F = int(shape[0]), int(shape[1]), int(max_frames)
        self.device = torch.device(device)
        self.lib = N.load()
        self.max_disparity, self.block = int(max_disparity), int(block)
        if not self.lib.svb_sgbm_supported(self.H, self.W, 128, self.block):
            raise ValueError("matcher: 128 disparities, odd block <= 31 and W - 128 > block / 2 are supported")
        fl = max(1, min(int(frames_in_flight), self.F))
        nb = max(self.lib.svb_sgbm_scratch_bytes(fl, self.H, self.W, 128), self.lib.svb_speckle_scratch_bytes(self.F, self.H, self.W))
        self.scratch = torch.empty((nb,), dtype=torch.uint8, device=self.device)
        self.disp16 = torch.empty((self.F, self.H, self.W), dtype=torch.int16, device=self.device)
        self.disp8 = torch.empty((self.F, self.H, self.W), dtype=torch.uint8, device=self.device)
        # bit 0 after compute(): a block cost exceeded 32767 - P2, where cv2's own output depends on its SIMD path
        self.status = torch.zeros((1,), dtype=torch.int32, device=self.device)

    @_on_device
    def compute(self, grayL, grayR):
        """stereoProcessor.compute for F pairs: int16 [F,H,W], disparity x 16, -16 = no match."""
        L, R = _as_cuda_u8(grayL, self.device), _as_cuda_u8(grayR, self.device)
        if L.dim() == 2:
            L, R = L[None], R[None]
        F = L.shape[0]
        if tuple(L.shape[1:]) != (self.H, self.W) or L.shape != R.shape or F > self.F:
            raise ValueError("expected up to %d pairs of %dx%d, got %s / %s" % (self.F, self.H, self.W, tuple(L.shape), tuple(R.shape)))
        self.status.zero_()
        N.check(self.lib.svb_sgbm_compute(N.ptr(L), N.ptr(R), F, self.H, self.W, 128, self.block, N.ptr(self.disp16),
                                          N.ptr(self.status), N.ptr(self.scratch), self.scratch.numel(), N.stream_ptr()),
                "svb_sgbm_compute")
        per = self.lib.svb_sgbm_scratch_bytes(1, self.H, self.W, 128) - 256           # bytes one frame in flight needs
        in_flight = max(1, min(F, (self.scratch.numel() - 256) // per))
        N.count(8 * ((F + in_flight - 1) // in_flight))      # per chunk: prefilter, hsum, vsum_v, d1, d3, h, final, median
        return self.disp16[:F]

    @_on_device
    def disparity(self, grayL, grayR):
        """functions.disparity for F pairs: uint8 [F,H,W] in half pixels, 0 = no match (crop: a view the caller takes)."""
        d16 = self.compute(grayL, grayR)
        F, st = d16.shape[0], N.stream_ptr()
        N.check(self.lib.svb_filter_speckles(N.ptr(d16), F, self.H, self.W, 0, 4000, self.max_disparity - 5, N.ptr(self.scratch),
                                             self.scratch.numel(), st), "svb_filter_speckles")
        N.check(self.lib.svb_disparity_scale(N.ptr(d16), F * self.H * self.W, self.max_disparity, N.ptr(self.disp8), st),
                "svb_disparity_scale")
        N.count(5)
        return self.disp8[:F]


class StereoPipeline:
    """Camera images to road mask for F chained frames of one sequence, all on the device: the batch form of
    stereovision.performStereoVision's stages 1-8 (stereovision.py:44-189) -- FrontEnd (gamma, gray, equalise) on both
    views, Matcher (disparity), RoadPath (fill, clouds, plane, road points), RoadCleaner (cleaning, hull, obstacles).
    The one-frame drop-in with the reference's signature is stereo.vision_b200/stereovision.py; this class is the
    throughput entry point (device-draw RANSAC, `seed`)."""

    def __init__(self, shape=(544, 1024), max_frames=16, frames_in_flight=None, trials=600, samples=600, step=2,
                 carmask=None, device="cuda:0"):
        F = int(max_frames)
        self.front = FrontEnd(shape, max_frames=F, device=device)
        self.front_r = FrontEnd(shape, max_frames=F, device=device)
        self.matcher = Matcher(shape, max_frames=F, frames_in_flight=frames_in_flight or F, device=device)
        self.path = RoadPath(shape, step=step, trials=trials, samples=samples, carmask=carmask, max_frames=F, device=device)
        self.cleaner = RoadCleaner(shape, max_frames=F, device=device)

    def process(self, bgrL, bgrR, prev=None, seed=0, obstacles=True):
        """bgrL, bgrR uint8 [F,H,W,3] -> dict(result = the path's BatchResult, disparity, cleaned, counts, points
        [, hull outputs]); result.filled[-1] is the `prev` of the next call."""
        colourL, grayL = self.front.process(bgrL)
        _, grayR = self.front_r.process(bgrR)
        disp = self.matcher.disparity(grayL, grayR)
        res = self.path.process(disp, colourL, prev=prev, seed=seed)
        cleaned, counts, points = self.cleaner.clean(res.road_mask)
        out = {"result": res, "disparity": disp, "cleaned": cleaned, "counts": counts, "points": points}
        if obstacles:
            out.update(self.cleaner.obstacles())
        return out

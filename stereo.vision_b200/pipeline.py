"""The fused, frame-batched path: stereovision.py:58-140 for F frames of one sequence per call.

`RoadPath.process` is the throughput entry point that sits beside the per-helper drop-in API
(functions.py of this package).  One call enqueues, on the current CUDA stream and without any host
synchronisation:

    svb_fill_sequence_counted    a1    temporal fill (the only cross-frame dependency, walked in-kernel)
                                       + the site counts of a4 (the bytes are in registers anyway)
    svb_project_compact_counted  a2-a4 scan / write: cloud (8 B records) + masked cloud (4 B)
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

KERNELS_PER_BATCH = 6      # fill+count, scan, write, score, finalize, road cluster kernel


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
        self.cloud = eng.cloud[:F]            # int64 [F,L] (uint64 records)
        self.mcloud = eng.mcloud[:F]          # int32 [F,L] (uint32 records)
        self.overlay = None

    def abc(self):
        """Host float64 [F,3]; NaN rows where no plane was found (RANSAC -> (None, None))."""
        return self.planes[:, :3].cpu().numpy()

    def best_trial(self):
        return self.planes[:, 6].cpu().numpy().astype(np.int64)


class RoadPath:
    def __init__(self, shape=(544, 1024), step=2, trials=600, samples=600, point_threshold=0.05,
                 colour_threshold=10, carmask=None, camera: Camera = DEFAULT_CAMERA, max_frames=1,
                 device="cuda:0", use_reproj_table=True):
        N.require_cuda()
        self.lib = N.load()
        self.device = torch.device(device)
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
        self.cloud = torch.empty((F, self.L), dtype=torch.int64, device=dev)
        self.mcloud = torch.empty((F, self.L), dtype=torch.int32, device=dev)
        self.counts = torch.zeros((F, 4), dtype=torch.int32, device=dev)
        self.errors = torch.empty((F, self.T), dtype=torch.float64, device=dev)
        self.planes = torch.zeros((F, N.PLANE_STRIDE), dtype=torch.float64, device=dev)
        self.hist = torch.zeros((F, N.HIST_STRIDE), dtype=torch.int32, device=dev)
        self.road_mask = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
        nb = max(self.lib.svb_project_scratch_bytes(F, H, W, self.step),
                 self.lib.svb_road_scratch_bytes(F, H, W, self.step))
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

    # ------------------------------------------------------------------
    def process(self, disp, bgr, prev=None, seed=0, frame0=0, indices=None, overlay=None,
                fill=True) -> BatchResult:
        """Run the path on disp uint8 [F,H,W] + bgr uint8 [F,H,W,3].

        prev     : filled disparity of the frame before disp[0] (None at the start of a sequence);
        indices  : (sample_idx int32 [F,T,S], triple_idx int32 [F,T,3]) -> parity mode (explicit
                   draws, e.g. the replay of Python's `random`); None -> counter-based on-device
                   draws keyed by (seed, frame0 + t, trial);
        overlay  : optional uint8 [F,H,W,3] CUDA tensor pre-filled with imgL; road sites get (0,255,0);
        fill     : False skips the temporal fill entirely (threshold_option other than 'previous').
        """
        lib, dev = self.lib, self.device
        disp = _as_cuda_u8(disp, dev)
        bgr = _as_cuda_u8(bgr, dev)
        if disp.dim() == 2:
            disp, bgr = disp[None], bgr[None]
        F = disp.shape[0]
        if F > self.Fmax or tuple(disp.shape[1:]) != (self.H, self.W) or tuple(bgr.shape) != (F, self.H, self.W, 3):
            raise ValueError("bad batch shape %s / %s" % (tuple(disp.shape), tuple(bgr.shape)))
        H, W, st = self.H, self.W, N.stream_ptr()
        res = BatchResult(self, F)
        if fill:
            # a1 + the counting pass of a4 in one kernel; the scratch carries the counts to the projection
            prev_t = _as_cuda_u8(prev, dev) if prev is not None else None
            N.check(lib.svb_fill_sequence_counted(N.ptr(disp), N.ptr(prev_t), N.ptr(self.filled), N.ptr(self.carmask),
                                                  F, H, W, self.step, N.ptr(self.scratch), self.scratch.numel(), st),
                    "svb_fill_sequence_counted")
            project = lib.svb_project_compact_counted
        else:
            self.filled[:F].copy_(disp)
            project = lib.svb_project_compact
        src = self.filled
        N.check(project(N.ptr(src), N.ptr(self.carmask), N.ptr(bgr), F, H, W, self.step,
                        C.byref(self._cam), N.ptr(self.cloud), N.ptr(self.mcloud), N.ptr(self.counts),
                        N.ptr(self.scratch), self.scratch.numel(), st), "svb_project_compact")
        si = ti = None
        if indices is not None:
            si = indices[0] if torch.is_tensor(indices[0]) else torch.from_numpy(np.ascontiguousarray(indices[0], dtype=np.int32))
            ti = indices[1] if torch.is_tensor(indices[1]) else torch.from_numpy(np.ascontiguousarray(indices[1], dtype=np.int32))
            si = si.to(dev).contiguous().view(F, self.T, self.S)
            ti = ti.to(dev).contiguous().view(F, self.T, 3)
        N.check(lib.svb_ransac_score(N.ptr(self.mcloud), 1, self.L, N.ptr(self.counts), 4, 1, F, self.T, self.S,
                                     self.max_m, C.byref(self._cam), N.ptr(si), N.ptr(ti), seed, frame0, N.ptr(self.errors), st),
                "svb_ransac_score")
        N.check(lib.svb_ransac_finalize(N.ptr(self.mcloud), 1, self.L, N.ptr(self.counts), 4, 1, F, self.T, self.S,
                                        C.byref(self._cam), N.ptr(ti), seed, frame0, N.ptr(self.errors),
                                        self.point_threshold, N.ptr(self.planes), st), "svb_ransac_finalize")
        ov = None
        if overlay is not None:
            ov = _as_cuda_u8(overlay, dev)
            res.overlay = ov
        N.check(lib.svb_road_points(N.ptr(self.cloud), N.ptr(self.counts), N.ptr(self.planes), F, H, W, self.step,
                                    C.byref(self._cam), self.colour_threshold, N.ptr(self.reproj_bits), N.ptr(self.hist), N.ptr(self.road_mask),
                                    N.ptr(ov), N.ptr(self.counts), N.ptr(self.scratch), self.scratch.numel(), st),
                "svb_road_points")
        N.count(KERNELS_PER_BATCH - 1 + self.road_launches + (0 if fill else 2))   # no fill: copy + count kernels
        return res

    def dump_indices(self, F, seed=0, frame0=0):
        """The counter-based draws the last process() call used (for the oracle): int32 [F,T,S], [F,T,3]."""
        si = torch.empty((F, self.T, self.S), dtype=torch.int32, device=self.device)
        ti = torch.empty((F, self.T, 3), dtype=torch.int32, device=self.device)
        N.check(self.lib.svb_ransac_dump_indices(N.ptr(self.mcloud), 1, self.L, N.ptr(self.counts), 4, 1, F, self.T,
                                                 self.S, C.byref(self._cam), seed, frame0, N.ptr(si), N.ptr(ti),
                                                 N.stream_ptr()), "svb_ransac_dump_indices")
        N.count(1)
        return si, ti

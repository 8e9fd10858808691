#!/usr/bin/env python
"""Micro-benchmark of the road-mask cleaning kernels (CUDA events, 300 resident frames of road masks
produced by the path itself on the synthetic sequence)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import stereo.vision_b200._native as N
import stereo.vision_b200.synth as synth
from stereo.vision_b200.pipeline import RoadCleaner, RoadPath

F, H, W = 300, 544, 1024
dev = torch.device("cuda:0")
eng = RoadPath((H, W), max_frames=F, device=dev)
disp, bgr = synth.make_sequence_device(0, F, dev) if hasattr(synth, "make_sequence_device") else (None, None)
if disp is None:
    fr = [synth.make_frame(0, t) for t in range(8)]
    disp = torch.from_numpy(np.stack([f[0] for f in fr])).to(dev).repeat((F + 7) // 8, 1, 1)[:F].contiguous()
    bgr = torch.from_numpy(np.stack([f[1] for f in fr])).to(dev).repeat((F + 7) // 8, 1, 1, 1)[:F].contiguous()
res = eng.process(disp, bgr, seed=1)
masks = res.road_mask if hasattr(res, "road_mask") else eng.road_mask[:F]
masks = masks.clone()
torch.cuda.synchronize()
print("road pixels per frame: %.0f" % float((masks != 0).sum() / F))


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


for bands, want_bytes in ((0, True), (4, True), (2, True), (4, False)):
    N.check(N.load().svb_debug_clean_bands(bands), "bands")
    cl = RoadCleaner((H, W), max_frames=F, device=dev, want_bytes=want_bytes)
    t_all = timeit(lambda: cl.clean(masks))
    t_clean = timeit(lambda: cl.clean(masks, points=False))
    torch.cuda.synchronize()
    n = int(cl.counts.sum().item())
    by = masks.numel() * (2 if want_bytes else 1) + cl.bits.numel() * 4
    print("bands %d bytes out=%s: clean %.1f us (%.0f GB/s algorithmic), clean+points %.1f us, %d points/frame, status %d"
          % (bands, want_bytes, t_clean, by / t_clean / 1e3, t_all, n // F, int(cl.status.item())))

#!/usr/bin/env python
"""Micro-benchmark of the obstacle stage (svb_road_hull) on the cleaned road masks of the synthetic sequence:
which of its outputs costs what (CUDA events, 300 resident frames)."""
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
fr = [synth.make_frame(0, t) for t in range(8)]
disp = torch.from_numpy(np.stack([f[0] for f in fr])).to(dev).repeat((F + 7) // 8, 1, 1)[:F].contiguous()
bgr = torch.from_numpy(np.stack([f[1] for f in fr])).to(dev).repeat((F + 7) // 8, 1, 1, 1)[:F].contiguous()
res = eng.process(disp, bgr, seed=1)
cl = RoadCleaner((H, W), max_frames=F, device=dev)
cl.clean(res.road_mask)
out = cl.obstacles()
torch.cuda.synchronize()
print("hull vertices per frame: mean %.1f max %d" % (float(out["hull_n"].float().mean()), int(out["hull_n"].max())))
lib = N.load()


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


sb = int(lib.svb_road_hull_scratch_bytes(F, H))
for name, hull, centre, masks in (("hull only", 1, 0, 0), ("hull + centre", 1, 1, 0), ("hull + masks", 1, 0, 1), ("all", 1, 1, 1),
                                  ("masks only", 0, 0, 1)):
    def run():
        N.check(lib.svb_road_hull(N.ptr(cl.bits), F, H, W, cl._hull_cap, N.ptr(cl.hull) if hull else None,
                                  N.ptr(cl.hull_n) if hull else None, N.ptr(cl.centre) if centre else None,
                                  N.ptr(cl.hull_mask) if masks else None, N.ptr(cl.obstacle_mask) if masks else None,
                                  N.ptr(cl._hull_scratch) if masks else None, sb if masks else 0, N.stream_ptr()), "svb_road_hull")
    print("%-14s %.1f us" % (name, timeit(run)))

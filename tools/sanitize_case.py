#!/usr/bin/env python
"""A small end-to-end case for compute-sanitizer (memcheck / racecheck / synccheck): two chained 1024x544 frames through the
fused path in every draw mode, the records road path, the cleaning kernels and the no-allocation compaction.

    compute-sanitizer --tool racecheck python tools/sanitize_case.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import stereo.vision_b200._native as N                      # noqa: E402
import stereo.vision_b200.synth as synth                    # noqa: E402
from stereo.vision_b200.masks import default_carmask        # noqa: E402
from stereo.vision_b200.pipeline import RoadCleaner, RoadPath   # noqa: E402

F, T, S = 2, 64, 400
disp, bgr = synth.make_sequence(3, F)
cm = default_carmask(disp.shape[1:], prefer_cwd=False)
for road_path in ("sites", "records"):
    eng = RoadPath(trials=T, samples=S, carmask=cm, max_frames=F, road_path=road_path)
    r = eng.process(disp, bgr, seed=11)
    r = eng.process(disp, bgr, py_seeds=[5, 6])
    torch.cuda.synchronize()
    print(road_path, r.counts.tolist())
cl = RoadCleaner(max_frames=F)
cleaned, counts, pts = cl.clean(r.road_mask)
cl.obstacles()
lib = N.load()
n = int(r.counts[0, 0])
dist = torch.rand((n,), dtype=torch.float64, device="cuda") * 0.1
out = torch.empty((n,), dtype=torch.int64, device="cuda")
cnt = torch.zeros((1,), dtype=torch.int32, device="cuda")
N.check(lib.svb_compact_cloud_static(N.ptr(eng.cloud[0]), n, 0, N.ptr(dist), None, 0.05, N.ptr(out), N.ptr(cnt), N.stream_ptr()), "static")
torch.cuda.synchronize()
print("ok", int(cnt), counts.tolist())

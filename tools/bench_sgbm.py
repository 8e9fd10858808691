#!/usr/bin/env python
"""Per-kernel times of the matcher stage on F synthetic 1024x544 pairs: tools/bench_sgbm.py [F] [in_flight]"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import stereo.vision_b200._native as N
import stereo.vision_b200.synth as synth
from stereo.vision_b200.pipeline import Matcher

F = int(sys.argv[1]) if len(sys.argv) > 1 else 8
fl = int(sys.argv[2]) if len(sys.argv) > 2 else F
H, W = 544, 1024
pairs = [synth.make_stereo_pair(1, t, (H, W)) for t in range(min(F, 4))]
L = torch.from_numpy(np.stack([pairs[t % len(pairs)][0] for t in range(F)])).cuda()
R = torch.from_numpy(np.stack([pairs[t % len(pairs)][1] for t in range(F)])).cuda()
m = Matcher(shape=(H, W), max_frames=F, frames_in_flight=fl)
for _ in range(2):
    m.disparity(L, R)
torch.cuda.synchronize()
N.profile_enable(True)
reps = 3
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    m.disparity(L, R)
e1.record()
torch.cuda.synchronize()
prof = N.profile_read()
N.profile_enable(False)
tot = e0.elapsed_time(e1) / reps
w1 = W - 128
vol = H * w1 * 128 * 2
alg = {"sgbm_prefilter": 2 * H * W * (1 + 16), "sgbm_hsum": 2 * H * W * 16 + vol, "sgbm_vsum": 2.25 * vol, "sgbm_path_v": 1.25 * vol,
       "sgbm_path_d1": 1.25 * vol, "sgbm_path_d3": 1.25 * vol, "sgbm_path_h": 1.25 * vol, "sgbm_final": 2 * vol + 2 * H * W,
       "sgbm_median": 4 * H * W, "speckle": H * W * (2 * 3 + 8 * 3), "disparity_scale": 3 * H * W}
print("F %d in flight %d: %.3f ms per call, %.3f ms per frame, %.0f frames/s" % (F, fl, tot, tot / F, F / tot * 1e3))
for k, (ms, n) in prof.items():
    per_call = ms / reps
    print("  %-16s %8.3f ms/call %7.1f us/frame  %6.0f GB/s (algorithmic)" % (k, per_call, per_call / F * 1e3, alg.get(k, 0) * F / per_call / 1e6))

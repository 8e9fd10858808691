#!/usr/bin/env python
"""Micro-benchmark of the fill entry points (CUDA events, 300 resident frames)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stereo.vision_b200._native as N
from stereo.vision_b200.masks import default_carmask

F, H, W = 300, 544, 1024
lib = N.load()
dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
raw = torch.from_numpy(rng.integers(0, 255, (F, H, W), dtype=np.uint8)).to(dev)
raw[torch.rand((F, H, W), device=dev) < 0.1] = 0
out = torch.empty_like(raw)
cm = torch.from_numpy(default_carmask((H, W), prefer_cwd=False)).to(dev)
nb = lib.svb_project_scratch_bytes(F, H, W, 2)
scratch = torch.empty((nb,), dtype=torch.uint8, device=dev)
st = N.stream_ptr()

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3

t_plain = timeit(lambda: N.check(lib.svb_fill_sequence(N.ptr(raw), None, N.ptr(out), F, H, W, st), "fill"))
t_cnt = timeit(lambda: N.check(lib.svb_fill_sequence_counted(N.ptr(raw), None, N.ptr(out), N.ptr(cm), F, H, W, 2, N.ptr(scratch), nb, st), "fillc"))
t_copy = timeit(lambda: out.copy_(raw))
print("fill plain %.1f us, fill+count %.1f us, torch copy %.1f us (%.0f GB/s)" % (t_plain, t_cnt, t_copy, 2 * raw.numel() / t_copy / 1e3))

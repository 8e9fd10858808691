#!/usr/bin/env python
"""One cleaning call on 300 synthetic road masks (the target of the ncu capture of the cleaning kernels)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from stereo.vision_b200.pipeline import RoadCleaner

F, H, W = 300, 544, 1024
dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
base = np.zeros((8, H, W), np.uint8)
for k in range(8):                                        # a road-like trapezoid of lattice points plus speckle
    for y in range(300, 540, 2):
        half = int((y - 280) * (1.2 + 0.05 * k))
        xs = np.arange(max(0, 500 - half), min(W, 500 + half), 2)
        base[k, y, xs[rng.random(len(xs)) < 0.8]] = 255
    base[k][rng.random((H, W)) < 0.002] = 255
masks = torch.from_numpy(base).to(dev).repeat(38, 1, 1)[:F].contiguous()
cl = RoadCleaner((H, W), max_frames=F, device=dev)
for _ in range(3):
    cl.clean(masks)
torch.cuda.synchronize()
print("points/frame", int(cl.counts.sum().item()) // F)

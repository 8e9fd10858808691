#!/usr/bin/env python
"""How far apart are the on-device draws and random.sample, and how far apart are two runs of random.sample?

    python tools/ks_spread.py            (on a GPU box; about 20 s)

Runs 240 frames (8 sequences x 30 chained frames, 600 trials x 400 samples) three times with CPython's draws replayed
on the device (three independent seed sets) and three times with the device draws (three seeds), and prints the
two-sample Kolmogorov-Smirnov p-values of the winner's error, K, R and the angle to the rendered plane for every
pairing, plus the medians.  ref-vs-ref is the calibration (0.3 - 0.99); ref-vs-dev must look the same.  This is the
instrument that showed the unscrambled congruential sample to be stratified (every dev median of K above every ref
median, p down to 0.007) and that chose between three scrambles (csrc/svb_ransac.cu, scramble()); the committed test
(tests/test_gpu_replay.py) runs a 2 x 2 subset of it."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from scipy.stats import ks_2samp

import stereo.vision_b200.synth as synth
from stereo.vision_b200.masks import default_carmask
from stereo.vision_b200.pipeline import RoadPath

NSEQ, F, T, S, RUNS = 8, 30, 600, 400, 3
COLS = ("best_error", "K", "R", "angle")


def observables(r, truth):
    pl, cn = r.planes.cpu().numpy(), r.counts.cpu().numpy()
    ang = np.degrees(np.arccos(np.clip(np.abs(((pl[:, :3] / pl[:, 3:4]) * truth).sum(axis=1)), 0, 1)))
    return np.stack([pl[:, 5], cn[:, 2].astype(float), cn[:, 3].astype(float), ang], axis=1)


def main():
    eng = RoadPath(trials=T, samples=S, carmask=default_carmask((544, 1024), prefer_cwd=False), max_frames=F)
    names = ["ref%d" % k for k in range(RUNS)] + ["dev%d" % k for k in range(RUNS)]
    acc = {n: [] for n in names}
    for seq in range(NSEQ):
        disp, bgr = synth.make_sequence(20 + seq, F)
        truth = np.stack([synth.true_plane(20 + seq, t) for t in range(F)])
        truth /= np.linalg.norm(truth, axis=1, keepdims=True)
        for k in range(RUNS):
            seeds = [synth.ransac_seed(20 + seq, t) + 7777 * k for t in range(F)]
            acc["ref%d" % k].append(observables(eng.process(disp, bgr, py_seeds=seeds), truth))
            acc["dev%d" % k].append(observables(eng.process(disp, bgr, seed=99 + seq + 1000 * k), truth))
    acc = {n: np.concatenate(v) for n, v in acc.items()}

    def row(a, b):
        return " ".join("%6.3f" % ks_2samp(acc[a][:, i], acc[b][:, i]).pvalue for i in range(4))
    print("KS p-values, columns:", " ".join(COLS))
    pairs = lambda xs: [(a, b) for i, a in enumerate(xs) for b in xs[i + 1:]]
    for a, b in pairs(names[:RUNS]):
        print("ref vs ref ", a, b, row(a, b))
    for a in names[:RUNS]:
        for b in names[RUNS:]:
            print("ref vs dev ", a, b, row(a, b))
    for a, b in pairs(names[RUNS:]):
        print("dev vs dev ", a, b, row(a, b))
    for n in names:
        print("median", n, " ".join("%10.4f" % np.median(acc[n][:, i]) for i in range(4)))


if __name__ == "__main__":
    main()

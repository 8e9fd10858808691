"""TEST INFRASTRUCTURE -- tests/golden/result_sha.json: shard.result_sha of sequences 0-7 (32 chained frames,
600 trials x 400 samples, draws = Python's `random` under synth.ransac_seed per frame) computed by the numpy
oracle on the CPU.  tests/test_gpu_configs.py::test_result_hash_is_the_single_gpu_hash holds the GPU to it.

    python oracle/make_result_sha.py
"""
import json
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
F, T, S = 32, 600, 400


def one(seq):
    import np_oracle as no
    import stereo.vision_b200.shard as shard
    import stereo.vision_b200.synth as synth
    from stereo.vision_b200.masks import default_carmask
    cm = default_carmask((544, 1024), prefer_cwd=False)
    masks, planes, counts = [], np.zeros((F, 8)), np.zeros((F, 4), np.int32)
    prev = None
    for t in range(F):
        d, c = synth.make_frame(seq, t)
        o = no.run_frame(d, c, prev, cm, seed=synth.ransac_seed(seq, t), trials=T, n_samples=S)
        prev = o["filled"]
        masks.append(o["road"]["mask"])
        planes[t, 6], planes[t, 7] = o["best"], 0.0 if o["abc"] is None else 1.0
        counts[t] = (len(o["cloud"]["d"]), len(o["mcloud"]["d"]), o["road"]["near_n"], o["road"]["road_n"])
    return str(seq), shard.result_sha(np.stack(masks), planes, counts)


if __name__ == "__main__":
    with mp.get_context("fork").Pool(8) as pool:
        sha = dict(pool.map(one, range(8)))
    out = {"frames": F, "trials": T, "samples": S, "made_by": "oracle/make_result_sha.py (numpy oracle, CPython random replay)", "sha": sha}
    with open(os.path.join(ROOT, "tests", "golden", "result_sha.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(out)

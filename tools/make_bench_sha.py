#!/usr/bin/env python
"""tests/golden/bench_sha.json: the result hashes bench.py prints per rank, computed on ONE GPU for sequences 0-7
with bench.py's own configuration (300 frames, 600 x 400, step 2, device draws keyed by 20260000 + rank).  A multi-GPU
run whose rank r does not reproduce sequence r's hash reports matches_single_gpu: false.

    gpurun -- 'python tools/make_bench_sha.py gpurun_out/bench_sha.json'   (then copy to tests/golden/)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                            # noqa: E402
import stereo.vision_b200.shard as shard                # noqa: E402

F, T, S, STEP = 300, 600, 400, 2
if __name__ == "__main__":
    import torch
    from stereo.vision_b200.masks import default_carmask
    from stereo.vision_b200.pipeline import RoadPath
    cm = default_carmask((bench.H, bench.W), prefer_cwd=False)
    eng = RoadPath((bench.H, bench.W), step=STEP, trials=T, samples=S, carmask=cm, max_frames=F, device="cuda:0")
    sha = {}
    for rank in range(8):
        disp, bgr = bench.make_sequence_parallel(shard.sequence_of_rank(rank), F, min(16, os.cpu_count() or 1))
        r = eng.process(torch.from_numpy(disp).cuda(), torch.from_numpy(bgr).cuda(), prev=None, seed=20260000 + rank, frame0=0)
        sha[str(shard.sequence_of_rank(rank))] = shard.result_sha(r.road_mask, r.planes, r.counts) + ":" + shard.planes_sha(r.planes)[:16]
        print(rank, sha[str(rank)], flush=True)
    out = {"frames": F, "trials": T, "samples": S, "step": STEP, "made_by": "tools/make_bench_sha.py on one B200", "sha": sha}
    json.dump(out, open(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "bench_sha.json"), "w"), indent=1)

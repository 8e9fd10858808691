"""world_size-2 gloo run of the N>1 host logic (rank -> sequence, timing reduction) on CPU."""
import json
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import hashlib, json, os, sys
    sys.path.insert(0, %r)
    import stereo.vision_b200.shard as shard
    import stereo.vision_b200.synth as synth
    rank, world = shard.init("gloo")
    seq = shard.sequence_of_rank(rank)
    d, c = synth.make_frame(seq, 0, (64, 128))
    ms = [10.0 + 5.0 * rank, 3.0 - rank]
    shard.barrier()
    mx = shard.max_over_ranks(ms)
    frames = shard.sum_over_ranks([7.0])
    import numpy as np
    planes = np.zeros((1, 8)); planes[0, 6] = rank; planes[0, 7] = 1.0
    mine = shard.result_sha((d > 40).astype(np.uint8) * 255, planes, np.full((1, 4), seq, np.int32))
    mine = mine + ":" + mine[:16]                     # the length bench.py sends (81 characters)
    everyone = shard.gather_strings(mine)
    print(json.dumps({"rank": rank, "world": world, "seq": seq, "sha": hashlib.sha256(d.tobytes()).hexdigest(),
                      "max": mx, "frames": frames, "result_sha": mine, "gathered": everyone}), flush=True)
    shard.finish()
""") % ROOT


def test_two_ranks_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541", WORLD_SIZE="2")
    procs = []
    for r in range(2):
        e = dict(env, RANK=str(r), LOCAL_RANK=str(r))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=e, stdout=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0
        outs.append(json.loads(out.strip().splitlines()[-1]))
    outs.sort(key=lambda o: o["rank"])
    assert [o["seq"] for o in outs] == [0, 1] and outs[0]["sha"] != outs[1]["sha"]
    for o in outs:
        assert o["world"] == 2 and o["max"] == [15.0, 3.0] and o["frames"] == [14.0]
        assert o["gathered"] == [outs[0]["result_sha"], outs[1]["result_sha"]]      # every rank sees every rank's hash
    assert len(outs[0]["result_sha"]) == 81 and outs[0]["result_sha"] != outs[1]["result_sha"]

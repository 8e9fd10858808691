"""TEST INFRASTRUCTURE -- builds the oracle's C restatement (oracle/sgbm_oracle.c) with gcc into
oracle/libsvb_oracle.so.  The .so is git-ignored but travels to the GPU box with the snapshot."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "libsvb_oracle.so")
SRC = [os.path.join(HERE, "sgbm_oracle.c")]


def build(force: bool = False) -> str:
    if not force and os.path.isfile(OUT) and all(os.path.getmtime(s) <= os.path.getmtime(OUT) for s in SRC):
        return OUT
    r = subprocess.run(["gcc", "-O2", "-fno-fast-math", "-shared", "-fPIC", "-o", OUT] + SRC, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("gcc failed building the oracle library")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))

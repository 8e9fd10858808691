"""Build libsvb200.so in-tree with nvcc for sm_100a (no JIT cache, the .so travels with the repo).

Every .cu is compiled to an object under csrc/build/ (in parallel) and the objects are linked into
stereo.vision_b200/libsvb200.so.  An object is reused only when the hash of its source, of every header
and of the flags is the one it was built from, so `build()` never trusts time stamps; `force=True`
(what __graft_entry__.build() asks for) recompiles everything.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
OUT = os.path.join(PKG, "libsvb200.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = ["svb_abi.cu", "svb_project.cu", "svb_ransac.cu", "svb_pyrandom.cu", "svb_road.cu", "svb_clean.cu", "svb_sites.cu",
           "svb_prep.cu", "svb_options.cu", "svb_hull.cu", "svb_sgbm.cu"]
HEADERS = [os.path.join(HERE, h) for h in ("svb_common.cuh", "svb_prof.cuh", "svb_geom.cuh")] + \
          [os.path.join(os.path.dirname(PKG), "include", h) for h in ("svb200.h", "svb200_debug.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-ffp-contract=off"]
LINK_LIBS = ["-ldl"]                   # nvtx3 is header-only and finds an attached tool through dlopen


def _digest(src: str) -> str:
    h = hashlib.sha256(" ".join(FLAGS).encode())
    for p in [src] + [x for x in HEADERS if os.path.isfile(x)]:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _stale():
    """sources whose object is missing or was built from other bytes"""
    out = []
    for s in SOURCES:
        obj = os.path.join(OBJ_DIR, s[:-3] + ".o")
        tag = obj + ".sha"
        want = _digest(os.path.join(HERE, s))
        if not (os.path.isfile(obj) and os.path.isfile(tag) and open(tag).read() == want):
            out.append((s, obj, tag, want))
    return out


def needs_build() -> bool:
    return not os.path.isfile(OUT) or bool(_stale())


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJ_DIR):
            os.remove(os.path.join(OBJ_DIR, f))
    todo = _stale()
    if not todo and os.path.isfile(OUT):
        return OUT
    flags = FLAGS + (["-Xptxas", "-v"] if verbose else [])

    def one(job):
        s, obj, tag, want = job
        r = subprocess.run([NVCC] + flags + ["-c", "-o", obj, os.path.join(HERE, s)], capture_output=True, text=True)
        if r.returncode == 0:
            with open(tag, "w") as f:
                f.write(want)
        return s, r

    with ThreadPoolExecutor(max_workers=min(len(todo), os.cpu_count() or 1) or 1) as ex:
        results = list(ex.map(one, todo))
    for s, r in results:
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("nvcc failed on %s" % s)
        if verbose:
            print(r.stdout + r.stderr)
    objs = [os.path.join(OBJ_DIR, s[:-3] + ".o") for s in SOURCES]
    r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs + LINK_LIBS,
                       capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("linking libsvb200.so failed")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))

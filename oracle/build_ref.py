"""TEST INFRASTRUCTURE -- the reference itself, compiled for the GPU box.

/root/reference is pure Python and exists only in the build container.  This recipe compiles its two modules
from the sources WHERE THEY LIE into byte code under oracle/_ref/ (git-ignored, shipped to the GPU box like the
built .so files) and copies the five mask PNGs the module reads at import time (functions.py:29-35):

    oracle/_ref/functions.code          /root/reference/functions.py, unmodified
    oracle/_ref/functions_s400.code     the same with ONE literal changed: random.sample(points, 600) -> 400
                                       (functions.py:286; BASELINE.json's configs quote 400 samples)
    oracle/_ref/stereovision.code       /root/reference/stereovision.py, unmodified
    oracle/_ref/masks/*.png

No source text of the reference is written anywhere.  Users: bench.py --impl reference / cpu_baseline (the timed
CPU arm, kind "reference") and tests/ (oracle/ref_compiled.py loads the modules).  The product never does.
"""
import importlib.util
import os
import py_compile
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = "/root/reference"


def build(force: bool = False) -> str:
    """returns oracle/_ref (built here when /root/reference is present, else whatever travelled with the repo)"""
    if not os.path.isfile(os.path.join(REF, "functions.py")):
        return OUT
    os.makedirs(os.path.join(OUT, "masks"), exist_ok=True)
    mode = py_compile.PycInvalidationMode.UNCHECKED_HASH
    for name in ("functions", "stereovision"):
        py_compile.compile(os.path.join(REF, name + ".py"), cfile=os.path.join(OUT, name + ".code"), dfile=name + ".py",
                           doraise=True, invalidation_mode=mode)
    src = open(os.path.join(REF, "functions.py")).read()
    lit = "random.sample(points, 600)"
    assert src.count(lit) == 1, "functions.py:286 changed"
    with tempfile.TemporaryDirectory() as tmp:                       # the edited text lives only in a temporary file
        p = os.path.join(tmp, "functions_s400.py")
        with open(p, "w") as fh:
            fh.write(src.replace(lit, "random.sample(points, 400)"))
        py_compile.compile(p, cfile=os.path.join(OUT, "functions_s400.code"), dfile="functions.py (sample literal 400)",
                           doraise=True, invalidation_mode=mode)
    for fn in os.listdir(os.path.join(REF, "masks")):
        if fn.endswith(".png"):
            shutil.copyfile(os.path.join(REF, "masks", fn), os.path.join(OUT, "masks", fn))
    with open(os.path.join(OUT, "BUILT_FROM"), "w") as fh:
        fh.write("%s  python %s  magic %s\n" % (REF, sys.version.split()[0], importlib.util.MAGIC_NUMBER.hex()))
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))

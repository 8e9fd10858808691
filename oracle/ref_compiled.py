"""TEST INFRASTRUCTURE -- loads the reference's own modules from the byte code oracle/build_ref.py left under
oracle/_ref/ (the reference tree itself is absent on the GPU box).  Used by bench.py's CPU arm and by tests/."""
import importlib.machinery
import importlib.util
import os
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_DIR, "functions.code")) and os.path.isdir(os.path.join(REF_DIR, "masks"))


def _load(pyc, modname):
    loader = importlib.machinery.SourcelessFileLoader(modname, os.path.join(REF_DIR, pyc))
    spec = importlib.util.spec_from_loader(modname, loader)
    mod = importlib.util.module_from_spec(spec)
    cwd = os.getcwd()
    os.chdir(REF_DIR)                       # the masks are read relative to the CWD at import time (functions.py:29-33)
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            loader.exec_module(mod)
    finally:
        os.chdir(cwd)
    return mod


def load_functions(n_samples: int = 600):
    """the reference's functions module (n_samples = 400: the build with functions.py:286's literal changed)"""
    if n_samples not in (600, 400):
        raise ValueError("compiled for 600 (unmodified) and 400 samples")
    return _load("functions.code" if n_samples == 600 else "functions_s400.code", "reference_functions_s%d" % n_samples)


def load_orchestrator(f):
    """the reference's stereovision module with `import functions as f` bound to `f` (stereovision.py:3)"""
    saved = sys.modules.get("functions")
    sys.modules["functions"] = f
    try:
        return _load("stereovision.code", "reference_stereovision_%d" % id(f))
    finally:
        if saved is None:
            sys.modules.pop("functions", None)
        else:
            sys.modules["functions"] = saved

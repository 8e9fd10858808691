"""CPU-side checks: the C-ABI library loads and exports every symbol include/svb200.h declares,
host-only entry points work, and the product refuses to run without a CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import np_oracle as no
import stereo.vision_b200._native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "svb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(svb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    lib = N.load()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), "libsvb200.so lacks %s" % n
        assert n in N.SIGNATURES, "binding lacks %s" % n
    assert sorted(N.SIGNATURES) == names
    assert lib.svb_version() == N.ABI_VERSION


def test_lattice_and_scratch_sizes():
    lib = N.load()
    for (H, W, step) in [(544, 1024, 2), (544, 1024, 1), (1088, 2048, 2), (37, 53, 3), (2, 2, 1)]:
        assert (lib.svb_lattice_rows(H, step), lib.svb_lattice_cols(W, step)) == no.lattice_shape(H, W, step)
    assert lib.svb_project_scratch_bytes(300, 544, 1024, 2) > 0
    assert lib.svb_road_scratch_bytes(1, 544, 1024, 2) >= 139264 // 8
    assert lib.svb_error_string(-4).decode().startswith("scratch")


def test_argument_errors_do_not_touch_the_gpu():
    lib = N.load()
    assert lib.svb_fill_sequence(None, None, None, 1, 4, 4, None) == -1
    assert lib.svb_project_compact(None, None, None, 1, 4, 4, 2, None, None, None, None, None, 0, None) == -1
    assert lib.svb_reproject(None, 3, 5, None, None, None, None) == -1


def test_plane_row_cutoff_is_exact():
    """cutoff = smallest v with fl(v/|abc|) >= thr, so |s-1| < cutoff <=> fl(|s-1|/|abc|) < thr."""
    lib = N.load()
    rng = np.random.default_rng(0)
    for _ in range(200):
        abc = rng.normal(size=3) * 10 ** rng.uniform(-2, 2)
        thr = float(10 ** rng.uniform(-3, 0))
        row = np.zeros(8)
        a = np.ascontiguousarray(abc)
        assert lib.svb_plane_from_abc(a.ctypes.data_as(C.c_void_p), thr, row.ctypes.data_as(C.c_void_p)) == 0
        nrm = np.sqrt(abc[0] * abc[0] + abc[1] * abc[1] + abc[2] * abc[2])
        assert row[3] == nrm and row[7] == 1.0
        v = row[4]
        assert v / nrm >= thr and np.nextafter(v, 0.0) / nrm < thr


def test_histogram_keys_match_reference_strings():
    from stereo.vision_b200.handles import bin_key
    import colorsys
    rng = np.random.default_rng(1)
    for r, g, b in rng.integers(0, 256, size=(3000, 3), dtype=np.uint8):
        key = str(round(colorsys.rgb_to_hsv(r, g, b)[0], 3))          # functions.py:76-78
        assert bin_key(int(no.hue_bin(r, g, b))) == key
    assert bin_key(0) == "0.0" and bin_key(1000) == "1.0"


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import stereo.vision_b200.functions as f
    from stereo.vision_b200.pipeline import RoadPath
    with pytest.raises(N.SvbError):
        RoadPath()
    with pytest.raises(N.SvbError):
        f.projectDisparityTo3d(np.zeros((8, 8), np.uint8), 128)
    from stereo.vision_b200.pipeline import Matcher
    import frame_harness as sv
    with pytest.raises(N.SvbError):
        Matcher()
    with pytest.raises(N.SvbError):
        f.disparity(np.zeros((40, 200), np.uint8), np.zeros((40, 200), np.uint8), 128, False)
    with pytest.raises(N.SvbError):                    # the first stage already needs the device; nothing is swallowed before it
        sv.performStereoVision(np.zeros((544, 1024, 3), np.uint8), np.zeros((544, 1024, 3), np.uint8), None, dict(sv.default_opts))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "stereo.vision_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(dp, fn)).read()
                assert "np_oracle" not in txt and "ref_harness" not in txt and "/root/reference" not in txt, fn
                assert "sgbm_oracle" not in txt and "libsvb_oracle" not in txt, fn

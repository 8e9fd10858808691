"""ctypes binding of libsvb200.so (include/svb200.h).

There is no CPU fallback: if the library is missing it is built with nvcc (csrc/build.py); if that
fails, or a call returns non-zero, an exception is raised.  Tensors are handed over as raw device
pointers; nothing here synchronises.
"""
import ctypes as C
import os

import torch

from .camera import Camera

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libsvb200.so")

HUE_BINS = 1001
HIST_STRIDE = 1024
PLANE_STRIDE = 8
HIST_SPANS = 1023            # SVB_HIST_SPANS: slot of a histogram row with the colour spans the sites kernel read
RANSAC_SHARE_SM = 1
ABI_VERSION = 15
SVB_OK, SVB_E_NULL, SVB_E_SHAPE, SVB_E_ARG, SVB_E_SCRATCH, SVB_E_LIMIT = 0, -1, -2, -3, -4, -5      # include/svb200.h
COMPACT_STATIC_MAX_N = 4194304
SVB_MAX_DIM = 2048                                                                  # csrc/svb_common.cuh


class SvbError(RuntimeError):
    pass


class _CCamera(C.Structure):
    _fields_ = [("focal_px", C.c_double), ("baseline_m", C.c_double),
                ("centre_h", C.c_double), ("centre_w", C.c_double)]


def ccam(cam: Camera) -> _CCamera:
    return _CCamera(cam.focal_px, cam.baseline_m, cam.centre_h, cam.centre_w)


P = C.c_void_p
I = C.c_int
L = C.c_int64
U64 = C.c_uint64
U32 = C.c_uint32
D = C.c_double
CAMP = C.POINTER(_CCamera)

# name -> (restype, argtypes); every symbol include/svb200.h declares
SIGNATURES = {
    "svb_version": (I, []),
    "svb_error_string": (C.c_char_p, [I]),
    "svb_lattice_rows": (I, [I, I]),
    "svb_lattice_cols": (I, [I, I]),
    "svb_fill_sequence": (I, [P, P, P, I, I, I, P]),
    "svb_fill_sequence_counted": (I, [P, P, P, P, I, I, I, I, P, L, P]),
    "svb_mask_disparity": (I, [P, P, P, I, I, I, P]),
    "svb_project_scratch_bytes": (L, [I, I, I, I]),
    "svb_project_compact": (I, [P, P, P, I, I, I, I, CAMP, P, P, P, P, L, P]),
    "svb_project_compact_counted": (I, [P, P, P, I, I, I, I, CAMP, P, P, P, P, L, P]),
    "svb_cloud_export": (I, [P, I, I, CAMP, P, P, P, P, P, P, P]),
    "svb_ransac_score": (I, [P, I, L, P, I, I, I, I, I, I, CAMP, P, P, U64, I, I, P, P]),
    "svb_ransac_dump_indices": (I, [P, I, L, P, I, I, I, I, I, CAMP, U64, I, P, P, P]),
    "svb_pyrandom_setsize": (L, [I]),
    "svb_ransac_replay_scratch_bytes": (L, [I, I, I]),
    "svb_ransac_replay_draws": (I, [P, I, L, P, I, I, I, I, I, I, CAMP, P, P, I, P, P, P, L, P]),
    "svb_ransac_finalize": (I, [P, I, L, P, I, I, I, I, I, CAMP, P, U64, I, P, D, P, P]),
    "svb_plane_from_abc": (I, [P, D, P]),
    "svb_point_errors": (I, [P, I, I, CAMP, P, P, P]),
    "svb_compact_scratch_bytes": (L, [I]),
    "svb_compact_cloud": (I, [P, I, I, P, P, D, P, P, P, L, P]),
    "svb_compact_cloud_static": (I, [P, I, I, P, P, D, P, P, P]),
    "svb_hue_histogram": (I, [P, I, P, P]),
    "svb_hue_bins": (I, [P, L, P, P]),
    "svb_reproject": (I, [P, I, I, CAMP, P, P, P]),
    "svb_reproject_table": (I, [I, I, CAMP, P, P]),
    "svb_scatter_points": (I, [P, I, P, I, I, I, P, P, P]),
    "svb_gamma_gray_hist": (I, [P, I, I, I, P, P, P, P, P]),
    "svb_equalize_lut": (I, [P, I, P, P]),
    "svb_apply_lut_frames": (I, [P, I, L, P, P, P]),
    "svb_lut_bytes": (I, [P, L, P, P, P]),
    "svb_fill_row_mean": (I, [P, I, I, I, P, P]),
    "svb_normal_line": (I, [P, P, I, P, P, I, I, I, CAMP, P, P, P]),
    "svb_road_hull": (I, [P, I, I, I, I, P, P, P, P, P, P, L, P]),
    "svb_road_hull_scratch_bytes": (L, [I, I]),
    "svb_sgbm_supported": (I, [I, I, I, I]),
    "svb_sgbm_scratch_bytes": (L, [I, I, I, I]),
    "svb_sgbm_compute": (I, [P, P, I, I, I, I, I, P, P, P, L, P]),
    "svb_speckle_scratch_bytes": (L, [I, I, I]),
    "svb_filter_speckles": (I, [P, I, I, I, I, I, I, P, L, P]),
    "svb_disparity_scale": (I, [P, L, I, P, P]),
    "svb_bits_words": (L, [I, I]),
    "svb_pack_mask": (I, [P, I, I, I, P, P, P]),
    "svb_clean_road": (I, [P, P, I, I, I, P, P, P, P, P]),
    "svb_clean_road_bits": (I, [P, P, I, I, I, P, P, P, P]),
    "svb_mask_points": (I, [P, P, I, I, I, I, P, P, P]),
    "svb_road_scratch_bytes": (L, [I, I, I, I]),
    "svb_road_points": (I, [P, P, P, I, I, I, I, CAMP, U32, P, P, P, P, P, P, L, P]),
    "svb_road_sites_supported": (I, [I, I, I]),
    "svb_road_sites_scratch_bytes": (L, [I, I, I, I]),
    "svb_road_points_sites": (I, [P, P, P, I, I, I, I, CAMP, U32, P, P, P, P, P, P, L, P]),
    "svb_reproject_bits_words": (L, [I, I]),
    "svb_reproject_bits": (I, [I, I, CAMP, P, P, P]),
    "svb_upload_rows": (I, [P, P, L, L, L, P]),
    "svb_upload_row_block": (I, [P, P, L, L, L, L, L, L, P]),
    "svb_profile_enable": (I, [I]),
    "svb_profile_kernels": (I, []),
    "svb_profile_get": (I, [I, C.c_char_p, I, C.POINTER(D), C.POINTER(I)]),
}

# include/svb200_debug.h: test / tuning hooks, not part of the boundary
DEBUG_SIGNATURES = {
    "svb_debug_force_two_pass": (I, [I]),
    "svb_debug_clean_bands": (I, [I]),
}

_lib = None


def load(build_if_missing: bool = True):
    """Load (building first if needed) the native library.  Raises if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing:
        from .csrc import build as _b
        try:
            _b.build()
        except Exception:
            if not os.path.isfile(LIB_PATH):
                raise
    if not os.path.isfile(LIB_PATH):
        raise SvbError("libsvb200.so is missing: run `python __graft_entry__.py build`")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in list(SIGNATURES.items()) + list(DEBUG_SIGNATURES.items()):
        fn = getattr(lib, name)           # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.svb_version() != ABI_VERSION:
        raise SvbError("libsvb200.so ABI %d != binding %d: rebuild" % (lib.svb_version(), ABI_VERSION))
    _lib = lib
    return lib


def check(code: int, what: str):
    if code != 0:
        msg = load().svb_error_string(code)
        raise SvbError("%s failed: %s (%d)" % (what, msg.decode() if msg else "?", code))


def require_cuda():
    if not torch.cuda.is_available():
        raise SvbError("stereo.vision_b200 needs a CUDA device (sm_100a); there is no CPU path")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise SvbError("expected a CUDA tensor")
    if not t.is_contiguous():
        raise SvbError("expected a contiguous tensor")
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# launch counter: every kernel launch the library performs on our behalf is accounted here
# (bench.py reports it as gpu_launches)
launches = 0


def count(n: int):
    global launches
    launches += n


def profile_enable(on: bool):
    check(load().svb_profile_enable(1 if on else 0), "svb_profile_enable")


def profile_read():
    """{kernel name: (total ms, launches)} since profile_enable(True); caller must have synchronised."""
    lib = load()
    out = {}
    for i in range(lib.svb_profile_kernels()):
        name = C.create_string_buffer(64)
        ms, n = D(0.0), I(0)
        check(lib.svb_profile_get(i, name, 64, C.byref(ms), C.byref(n)), "svb_profile_get")
        if n.value:
            out[name.value.decode()] = (ms.value, n.value)
    return out

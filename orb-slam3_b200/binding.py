"""ctypes binding of liborbx.so (the C ABI declared in include/orbx.h).

The library is built in-tree by build.py (nvcc, sm_100a) and MUST be present: there is no CPU or
PyTorch fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import pathlib

import numpy as np

_PKG = pathlib.Path(__file__).resolve().parent
LIB_PATH = _PKG / "liborbx.so"

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

ORBX_EMPTY_IMAGE, ORBX_ERR_CAPACITY, ORBX_ERR_GEOMETRY, ORBX_ERR_ARG, ORBX_ERR_CUDA, ORBX_ERR_STATE = -1, -2, -3, -4, -5, -6

# every symbol include/orbx.h declares (checked by tests/test_abi.py against the header)
SYMBOLS = {
    "orbx_last_error": (C.c_char_p, []),
    "orbx_version": (C.c_int, []),
    "orbx_device_count": (C.c_int, []),
    "orbx_create": (C.c_int, [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "orbx_destroy": (None, [C.c_void_p]),
    "orbx_levels": (C.c_int, [C.c_void_p]),
    "orbx_scale_factor": (C.c_float, [C.c_void_p]),
    "orbx_scale_factors": (C.POINTER(C.c_float), [C.c_void_p]),
    "orbx_inv_scale_factors": (C.POINTER(C.c_float), [C.c_void_p]),
    "orbx_level_sigma2": (C.POINTER(C.c_float), [C.c_void_p]),
    "orbx_inv_level_sigma2": (C.POINTER(C.c_float), [C.c_void_p]),
    "orbx_features_per_level": (C.POINTER(C.c_int), [C.c_void_p]),
    "orbx_extract": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "orbx_extract_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
                                     C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_set_rectify_maps": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t]),
    "orbx_prepare_extract": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                       C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "orbx_prepare_extract_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
                                             C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_prepare_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t]),
    "orbx_prepare_resident": (C.c_int, [C.c_void_p]),
    "orbx_get_prepared": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "orbx_max_keypoints": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "orbx_reserve": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "orbx_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t]),
    "orbx_run_resident": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "orbx_sync": (C.c_int, [C.c_void_p]),
    "orbx_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_get_pyramid_level": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
                                         C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "orbx_debug_level_keys": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]),
    "orbx_debug_blurred_level": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]),
    "orbx_last_stage_ms": (C.c_float, [C.c_void_p, C.c_int]),
    "orbx_set_profiling": (None, [C.c_void_p, C.c_int]),
    "orbx_timer_start": (C.c_int, [C.c_void_p]),
    "orbx_timer_stop": (C.c_float, [C.c_void_p]),
    "orbx_debug_fast_score_map": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_void_p]),
    "orbx_launch_count": (C.c_longlong, [C.c_void_p]),
    "orbx_hamming256": (C.c_int, [C.c_void_p, C.c_void_p]),
    "orbx_stereo_match": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int]),
    "orbx_stereo_match_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int]),
    "orbx_stereo_match_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float]),
    "orbx_extract_stereo": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                      C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "orbx_knn2_hamming": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_search_by_projection_frame_rig": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p] + [C.c_float] * 4 +
                                            [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 10 + [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_search_by_bow_rig": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_float, C.c_int, C.c_void_p]),
    "orbx_search_for_triangulation_cb": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                   C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                   C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "orbx_stereo_fisheye_match": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "orbx_extract_batch_multi": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_int, C.c_int,
                                           C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_knn2_hamming_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "orbx_search_by_projection": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                            C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int,
                                            C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_float, C.c_float, C.c_int, C.c_float, C.c_void_p]),
    "orbx_window_candidates": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_float, C.c_float, C.c_float, C.c_float,
                                         C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "orbx_search_by_projection_frame": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                                  C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int,
                                                  C.c_int] + [C.c_void_p] * 8 +
                                        [C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_search_by_projection_keyframe": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                     C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int,
                                                     C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "orbx_search_for_initialization": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                 C.c_float, C.c_float, C.c_float, C.c_float,
                                                 C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]),
    "orbx_search_for_triangulation": (C.c_int, [C.c_int] + ([C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int]) * 2 +
                                      [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p] + [C.c_int] * 4 + [C.c_void_p]),
    "orbx_search_by_bow": (C.c_int, [C.c_int, C.c_int] + ([C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_int]) * 2 +
                           [C.c_float, C.c_int, C.c_void_p]),
    "orbx_vocabulary_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.POINTER(C.c_void_p)]),
    "orbx_vocabulary_destroy": (None, [C.c_void_p]),
    "orbx_bow_transform": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 3),
    "orbx_compute_bow": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_int),
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]),
    "orbx_distinctive_descriptors": (C.c_int, [C.c_int, C.c_int] + [C.c_void_p] * 4),
    "orbx_search_by_projection_rig": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p] +
                                      [C.c_float] * 4 + [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 12 +
                                      [C.c_float, C.c_float, C.c_int, C.c_float, C.c_void_p]),
    "orbx_undistort_keypoints": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int, C.c_void_p]),
    "orbx_search_local_points": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p] + [C.c_float] * 4 + [C.c_void_p, C.c_int] +
                                 [C.c_void_p] * 3 + [C.c_float] * 6 + [C.c_int] + [C.c_void_p] * 6 + [C.c_float] * 3 + [C.c_int, C.c_float] +
                                 [C.c_void_p] * 8),
    "orbx_search_by_projection_sim3": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                                 C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "orbx_fuse": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float,
                            C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_void_p, C.c_void_p]),
    "orbx_search_by_sim3": (C.c_int, [C.c_int] + [C.c_void_p, C.c_void_p, C.c_int] * 2 + [C.c_float] * 4 + [C.c_void_p, C.c_int] +
                            [C.c_void_p] * 10 + [C.c_float, C.c_void_p]),
    "orbx_three_maxima": (None, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load liborbx.so; raises (never falls back) when the CUDA library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI drifted
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class OrbxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"orbx error {code}: {msg}")
        self.code = code


def check(code: int) -> int:
    if code < 0:
        raise OrbxError(code, load().orbx_last_error().decode(errors="replace"))
    return code

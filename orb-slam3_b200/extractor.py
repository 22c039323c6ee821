"""Python mirror of ORB_SLAM3::ORBextractor (reference include/ORBextractor.h:49-83) on the C ABI.

Only used by tests / bench / smoke; the production host side is the C++ shim in cpp/.  Every call
goes through liborbx.so (CUDA); nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import binding
from .binding import KP_DTYPE, check


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) — src/ORBextractor.cc:468."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, device=0):
        self.lib = binding.load()
        h = C.c_void_p()
        check(self.lib.orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, C.byref(h)))
        self.h = h
        self.nfeatures, self.nlevels, self.device = nfeatures, nlevels, device

    def close(self):
        if getattr(self, "h", None):
            self.lib.orbx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- getters (include/ORBextractor.h:61-81) ----
    def _arr(self, fn, dtype):
        return np.ctypeslib.as_array(fn(self.h), shape=(self.nlevels,)).astype(dtype, copy=True)

    def GetLevels(self): return self.lib.orbx_levels(self.h)
    def GetScaleFactor(self): return self.lib.orbx_scale_factor(self.h)
    def GetScaleFactors(self): return self._arr(self.lib.orbx_scale_factors, np.float32)
    def GetInverseScaleFactors(self): return self._arr(self.lib.orbx_inv_scale_factors, np.float32)
    def GetScaleSigmaSquares(self): return self._arr(self.lib.orbx_level_sigma2, np.float32)
    def GetInverseScaleSigmaSquares(self): return self._arr(self.lib.orbx_inv_level_sigma2, np.float32)
    def features_per_level(self): return self._arr(self.lib.orbx_features_per_level, np.int32)

    def max_keypoints(self, rows, cols):
        return check(self.lib.orbx_max_keypoints(self.h, rows, cols))

    # ---- operator() (include/ORBextractor.h:57-59) ----
    def __call__(self, image: np.ndarray, lapping=(0, 0)):
        """Returns (monoIndex, keypoints, descriptors); monoIndex == -1 for an empty image."""
        if image is None or image.size == 0:
            kp = np.empty(0, KP_DTYPE); d = np.empty((0, 32), np.uint8); n = C.c_int(0)
            r = self.lib.orbx_extract(self.h, None, 0, 0, 0, int(lapping[0]), int(lapping[1]), kp.ctypes.data, d.ctypes.data, 0, C.byref(n))
            return r, kp, d
        assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1, "CV_8UC1 expected"
        cap = self.max_keypoints(*image.shape)
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        n = C.c_int(0)
        self._last_shape = (1,) + tuple(image.shape)
        mono = check(self.lib.orbx_extract(self.h, image.ctypes.data, image.shape[0], image.shape[1], image.strides[0],
                                           int(lapping[0]), int(lapping[1]), kps.ctypes.data, desc.ctypes.data, cap, C.byref(n)))
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, frames: np.ndarray, lapping=(0, 0), kps=None, desc=None):
        """frames: [B, rows, cols] uint8 (C-contiguous).  Returns (mono[B], nkp[B], kps[B,cap], desc[B,cap,32])."""
        assert frames.dtype == np.uint8 and frames.ndim == 3 and frames.flags.c_contiguous
        B, rows, cols = frames.shape
        cap = self.max_keypoints(rows, cols)
        if kps is None:
            kps = np.empty((B, cap), KP_DTYPE)
        if desc is None:
            desc = np.empty((B, cap, 32), np.uint8)
        nkp = np.empty(B, np.int32); mono = np.empty(B, np.int32)
        self._last_shape = (B, rows, cols)
        check(self.lib.orbx_extract_batch(self.h, B, frames.ctypes.data, rows, cols, frames.strides[1], frames.strides[0],
                                          int(lapping[0]), int(lapping[1]), kps.ctypes.data, desc.ctypes.data, cap,
                                          nkp.ctypes.data, mono.ctypes.data))
        return mono, nkp, kps, desc

    # ---- frame preparation in front of operator(): remap + cvtColor on the device ----
    def set_rectify_maps(self, mapx, mapy):
        """M1, M2 of cv::initUndistortRectifyMap(..., CV_32F, ...) (Examples/Stereo/stereo_euroc.cc:117-118); None clears."""
        if mapx is None:
            check(self.lib.orbx_set_rectify_maps(self.h, None, None, 0, 0, 0)); self._map_shape = None
            return
        mapx = np.ascontiguousarray(mapx, np.float32); mapy = np.ascontiguousarray(mapy, np.float32)
        assert mapx.shape == mapy.shape and mapx.ndim == 2
        check(self.lib.orbx_set_rectify_maps(self.h, mapx.ctypes.data, mapy.ctypes.data, mapx.shape[0], mapx.shape[1], mapx.strides[0]))
        self._map_shape = mapx.shape

    def _prepared_shape(self, src_shape):
        return getattr(self, "_map_shape", None) or tuple(src_shape[:2])

    def prepare_extract(self, image: np.ndarray, lapping=(0, 0), rgb=False):
        """cv::remap (stereo_euroc.cc:168) -> cvtColor (Tracking.cc:1264-1290) -> operator(); image HxW or HxWxC u8."""
        assert image.dtype == np.uint8 and image.ndim in (2, 3) and image.flags.c_contiguous
        ch = 1 if image.ndim == 2 else image.shape[2]
        rows, cols = self._prepared_shape(image.shape)
        cap = self.max_keypoints(rows, cols)
        kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
        n = C.c_int(0)
        self._last_shape = (1, rows, cols)
        mono = check(self.lib.orbx_prepare_extract(self.h, image.ctypes.data, image.shape[0], image.shape[1], ch, int(rgb),
                                                   image.strides[0], int(lapping[0]), int(lapping[1]),
                                                   kps.ctypes.data, desc.ctypes.data, cap, C.byref(n)))
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def prepare_extract_batch(self, frames: np.ndarray, lapping=(0, 0), rgb=False, kps=None, desc=None):
        """frames: [B, rows, cols] or [B, rows, cols, C] uint8.  Returns (mono[B], nkp[B], kps[B,cap], desc[B,cap,32])."""
        assert frames.dtype == np.uint8 and frames.ndim in (3, 4) and frames.flags.c_contiguous
        B = frames.shape[0]; ch = 1 if frames.ndim == 3 else frames.shape[3]
        rows, cols = self._prepared_shape(frames.shape[1:])
        cap = self.max_keypoints(rows, cols)
        if kps is None:
            kps = np.empty((B, cap), KP_DTYPE)
        if desc is None:
            desc = np.empty((B, cap, 32), np.uint8)
        nkp = np.empty(B, np.int32); mono = np.empty(B, np.int32)
        self._last_shape = (B, rows, cols)
        check(self.lib.orbx_prepare_extract_batch(self.h, B, frames.ctypes.data, frames.shape[1], frames.shape[2], ch, int(rgb),
                                                  frames.strides[1], frames.strides[0], int(lapping[0]), int(lapping[1]),
                                                  kps.ctypes.data, desc.ctypes.data, cap, nkp.ctypes.data, mono.ctypes.data))
        return mono, nkp, kps, desc

    def prepare_upload(self, frames: np.ndarray, rgb=False):
        assert frames.dtype == np.uint8 and frames.ndim in (3, 4) and frames.flags.c_contiguous
        B = frames.shape[0]; ch = 1 if frames.ndim == 3 else frames.shape[3]
        rows, cols = self._prepared_shape(frames.shape[1:])
        check(self.lib.orbx_prepare_upload(self.h, B, frames.ctypes.data, frames.shape[1], frames.shape[2], ch, int(rgb),
                                           frames.strides[1], frames.strides[0]))
        self._shape = self._last_shape = (B, rows, cols)

    def prepare_resident(self): check(self.lib.orbx_prepare_resident(self.h))

    def prepared(self, frame=0) -> np.ndarray:
        """imRect / mImGray: the frame operator() saw."""
        _, rows, cols = self._last_shape
        out = np.empty((rows, cols), np.uint8)
        check(self.lib.orbx_get_prepared(self.h, frame, out.ctypes.data, out.strides[0]))
        return out

    # ---- device-resident path (kernel timing) ----
    def reserve(self, rows, cols, batch): check(self.lib.orbx_reserve(self.h, rows, cols, batch))

    def upload(self, frames: np.ndarray):
        assert frames.dtype == np.uint8 and frames.ndim == 3 and frames.flags.c_contiguous
        B, rows, cols = frames.shape
        check(self.lib.orbx_upload(self.h, B, frames.ctypes.data, rows, cols, frames.strides[1], frames.strides[0]))
        self._shape = self._last_shape = (B, rows, cols)

    def run_resident(self, lapping=(0, 0)): check(self.lib.orbx_run_resident(self.h, int(lapping[0]), int(lapping[1])))
    def sync(self): check(self.lib.orbx_sync(self.h))

    def download(self):
        B, rows, cols = self._shape
        cap = self.max_keypoints(rows, cols)
        kps = np.empty((B, cap), KP_DTYPE); desc = np.empty((B, cap, 32), np.uint8)
        nkp = np.empty(B, np.int32); mono = np.empty(B, np.int32)
        check(self.lib.orbx_download(self.h, kps.ctypes.data, desc.ctypes.data, cap, nkp.ctypes.data, mono.ctypes.data))
        return mono, nkp, kps, desc

    def timer_start(self): check(self.lib.orbx_timer_start(self.h))
    def timer_stop(self): return float(self.lib.orbx_timer_stop(self.h))
    def set_profiling(self, on=True): self.lib.orbx_set_profiling(self.h, int(on))
    def stage_ms(self, stage=-1): return float(self.lib.orbx_last_stage_ms(self.h, stage))
    def launch_count(self): return int(self.lib.orbx_launch_count(self.h))

    # ---- mvImagePyramid and stage dumps ----
    def pyramid_level(self, level, frame=0, padded=False) -> np.ndarray:
        r = C.c_int(); c = C.c_int()
        check(self.lib.orbx_get_pyramid_level(self.h, frame, level, int(padded), None, 0, C.byref(r), C.byref(c)))
        out = np.empty((r.value, c.value), np.uint8)
        check(self.lib.orbx_get_pyramid_level(self.h, frame, level, int(padded), out.ctypes.data, out.strides[0], C.byref(r), C.byref(c)))
        return out

    def blurred_level(self, level, frame=0) -> np.ndarray:
        r = C.c_int(); c = C.c_int()
        check(self.lib.orbx_get_pyramid_level(self.h, frame, level, 0, None, 0, C.byref(r), C.byref(c)))
        out = np.empty((r.value, c.value), np.uint8)
        check(self.lib.orbx_debug_blurred_level(self.h, frame, level, out.ctypes.data, out.strides[0]))
        return out

    def level_keys(self, level, stage, frame=0):
        """stage 0: FAST candidates (x, y, score) in reference order; 1: quadtree selection (+ angles)."""
        n = check(self.lib.orbx_debug_level_keys(self.h, frame, level, stage, None, None, 0))
        out = np.empty((n, 3), np.int32); ang = np.empty(n, np.float32)
        check(self.lib.orbx_debug_level_keys(self.h, frame, level, stage, out.ctypes.data, ang.ctypes.data, n))
        return (out, ang) if stage == 1 else out


def DescriptorDistance(a: np.ndarray, b: np.ndarray) -> int:
    """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2911-2931)."""
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return binding.load().orbx_hamming256(a.ctypes.data, b.ctypes.data)


def extract_batch_multi(extractors, frames: np.ndarray, lapping=(0, 0)):
    """One host batch over several extractor handles / devices (orbx_extract_batch_multi): contiguous frame ranges, one
    host thread per handle, results in frame order.  Returns (mono[B], nkp[B], kps[B,cap], desc[B,cap,32])."""
    assert frames.dtype == np.uint8 and frames.ndim == 3 and frames.flags.c_contiguous and len(extractors) > 0
    B, rows, cols = frames.shape
    cap = extractors[0].max_keypoints(rows, cols)
    kps = np.empty((B, cap), KP_DTYPE); desc = np.empty((B, cap, 32), np.uint8)
    nkp = np.empty(B, np.int32); mono = np.empty(B, np.int32)
    hs = (C.c_void_p * len(extractors))(*[x.h for x in extractors])
    check(binding.load().orbx_extract_batch_multi(hs, len(extractors), B, frames.ctypes.data, rows, cols, frames.strides[1], frames.strides[0],
                                                  int(lapping[0]), int(lapping[1]), kps.ctypes.data, desc.ctypes.data, cap,
                                                  nkp.ctypes.data, mono.ctypes.data))
    per = B // len(extractors); rem = B % len(extractors)
    for d, x in enumerate(extractors):
        n = per + (1 if d < rem else 0)
        if n:
            x._last_shape = (n, rows, cols)
    return mono, nkp, kps, desc

"""ctypes binding of the CPU ORACLE (oracle/liborc.so).  TEST INFRASTRUCTURE ONLY.

May be imported from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs — never from the product package.
"""
from __future__ import annotations

import ctypes as C
import pathlib
import subprocess

import numpy as np

_DIR = pathlib.Path(__file__).resolve().parent

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
CAND_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "<i4")])
assert KP_DTYPE.itemsize == 28 and CAND_DTYPE.itemsize == 12

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)


def build(force: bool = False) -> None:
    """Compile the C restatement (gcc, a second or two)."""
    if force:
        subprocess.run(["make", "-C", str(_DIR), "clean"], check=True, capture_output=True)
    subprocess.run(["make", "-C", str(_DIR)], check=True, capture_output=True)


def _ptr(a, t):
    return a.ctypes.data_as(t)


class Oracle:
    def __init__(self, fma: bool = False):
        path = _DIR / "_build" / ("liborc_fma.so" if fma else "liborc.so")
        if not path.exists():
            build()
        L = self.L = C.CDLL(str(path))
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_set_modes.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_extract.restype = C.c_int
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_int, i32p]
        for name in ("orc_scale_factors", "orc_inv_scale_factors", "orc_level_sigma2", "orc_inv_level_sigma2"):
            getattr(L, name).restype = f32p
            getattr(L, name).argtypes = [C.c_void_p]
        for name in ("orc_features_per_level", "orc_umax"):
            getattr(L, name).restype = i32p
            getattr(L, name).argtypes = [C.c_void_p]
        L.orc_remap_linear_u8.restype = None
        L.orc_remap_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_size_t]
        L.orc_undistort_points.restype = None
        L.orc_undistort_points.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_cvt_gray_u8.restype = None
        L.orc_cvt_gray_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.orc_levels.argtypes = [C.c_void_p]
        L.orc_level_dims.argtypes = [C.c_void_p, C.c_int, i32p, i32p]
        L.orc_level_plane.restype = C.c_void_p
        L.orc_level_plane.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
        L.orc_level_blurred.restype = C.c_void_p
        L.orc_level_blurred.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
        L.orc_level_candidates.restype = C.c_int
        L.orc_level_candidates.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.orc_level_selected.restype = C.c_int
        L.orc_level_selected.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_sinf.restype = C.c_float
        L.orc_sinf.argtypes = [C.c_float]
        L.orc_cosf.restype = C.c_float
        L.orc_cosf.argtypes = [C.c_float]
        L.orc_cvround_f.argtypes = [C.c_float]
        L.orc_resize_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_size_t]
        L.orc_border_reflect101.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int]
        L.orc_fast9_16.restype = C.c_int
        L.orc_fast9_16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_gaussian7x7_s2.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_size_t]
        L.orc_hamming256.restype = C.c_int
        L.orc_hamming256.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_bf_knn2.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_distribute_octree.restype = C.c_int
        L.orc_distribute_octree.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_ic_angle.restype = C.c_float
        L.orc_ic_angle.argtypes = [C.c_void_p, C.c_size_t, i32p]
        L.orc_orb_descriptor.argtypes = [C.c_void_p, C.c_size_t, C.c_float, C.c_int, C.c_void_p]
        L.orc_stereo_match.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_grid_build.restype = C.c_void_p
        L.orc_grid_build.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]
        L.orc_grid_destroy.argtypes = [C.c_void_p]
        L.orc_grid_query.restype = C.c_int
        L.orc_grid_query.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_search_by_projection.restype = C.c_int
        L.orc_search_by_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                               C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_float, C.c_float, C.c_int, C.c_float, C.c_void_p]
        L.orc_three_maxima.restype = None
        L.orc_three_maxima.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_search_by_projection_frame.restype = C.c_int
        L.orc_search_by_projection_frame.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int] + \
            [C.c_void_p] * 8 + [C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_search_for_initialization.restype = C.c_int
        L.orc_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]
        L.orc_distinctive_descriptors.restype = None
        L.orc_distinctive_descriptors.argtypes = [C.c_int] + [C.c_void_p] * 4
        L.orc_bow_transform.restype = None
        L.orc_bow_transform.argtypes = [C.c_void_p] * 5 + [C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4
        L.orc_bow_vectors.restype = C.c_int
        L.orc_bow_vectors.argtypes = [C.c_int] + [C.c_void_p] * 8 + [C.POINTER(C.c_int)]
        L.orc_search_by_projection_rig.restype = C.c_int
        L.orc_search_by_projection_rig.argtypes = [C.c_void_p, C.c_void_p, C.c_int] * 2 + [C.c_void_p] * 4 + [C.c_int] + [C.c_void_p] * 12 + \
            [C.c_float, C.c_float, C.c_int, C.c_float, C.c_void_p]
        L.orc_is_in_frustum.restype = C.c_int
        L.orc_is_in_frustum.argtypes = [C.c_void_p] * 3 + [C.c_float] * 10 + [C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_float] + [C.c_void_p] * 7
        L.orc_search_by_projection_sim3.restype = C.c_int
        L.orc_search_by_projection_sim3.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 5 + \
            [C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_fuse.restype = C.c_int
        L.orc_fuse.argtypes = [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_void_p, C.c_void_p]
        L.orc_search_by_sim3.restype = C.c_int
        L.orc_search_by_sim3.argtypes = ([C.c_void_p] * 3 + [C.c_int]) * 2 + [C.c_void_p] + [C.c_void_p] * 10 + [C.c_float, C.c_void_p]
        L.orc_search_by_projection_keyframe.restype = C.c_int
        L.orc_search_by_projection_keyframe.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 6 + \
            [C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_search_by_bow.restype = C.c_int
        L.orc_search_by_bow.argtypes = [C.c_int] + ([C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_int]) * 2 + \
            [C.c_float, C.c_int, C.c_void_p]
        L.orc_search_for_triangulation.restype = C.c_int
        L.orc_search_for_triangulation.argtypes = ([C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int]) * 2 + \
            [C.c_void_p] * 4 + [C.c_int] * 4 + [C.c_void_p]

    # ---- primitives ----
    def resize(self, src: np.ndarray, drows: int, dcols: int) -> np.ndarray:
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.empty((drows, dcols), np.uint8)
        self.L.orc_resize_linear_u8(src.ctypes.data, src.shape[0], src.shape[1], src.strides[0],
                                    dst.ctypes.data, drows, dcols, dst.strides[0])
        return dst

    def border101(self, img: np.ndarray, border: int) -> np.ndarray:
        h, w = img.shape
        out = np.zeros((h + 2 * border, w + 2 * border), np.uint8)
        out[border:border + h, border:border + w] = img
        self.L.orc_border_reflect101(out.ctypes.data, h, w, out.strides[0], border)
        return out

    def fast(self, img: np.ndarray, threshold: int, nonmax: bool = True) -> np.ndarray:
        img = np.ascontiguousarray(img, np.uint8)
        cap = img.size
        out = np.empty(cap, CAND_DTYPE)
        n = self.L.orc_fast9_16(img.ctypes.data, img.shape[0], img.shape[1], img.strides[0], threshold,
                                int(nonmax), out.ctypes.data, cap)
        return out[:n].copy()

    def gaussian(self, img: np.ndarray) -> np.ndarray:
        img = np.ascontiguousarray(img, np.uint8)
        dst = np.empty_like(img)
        self.L.orc_gaussian7x7_s2(img.ctypes.data, img.shape[0], img.shape[1], img.strides[0], dst.ctypes.data, dst.strides[0])
        return dst

    def remap(self, src: np.ndarray, mapx: np.ndarray, mapy: np.ndarray) -> np.ndarray:
        """cv2.remap(src, mapx, mapy, INTER_LINEAR); src HxW or HxWxC u8, maps float32 of the output size."""
        src = np.ascontiguousarray(src, np.uint8)
        ch = 1 if src.ndim == 2 else src.shape[2]
        mapx = np.ascontiguousarray(mapx, np.float32); mapy = np.ascontiguousarray(mapy, np.float32)
        dr, dc = mapx.shape
        dst = np.empty((dr, dc) if src.ndim == 2 else (dr, dc, ch), np.uint8)
        self.L.orc_remap_linear_u8(_ptr(src, C.c_void_p), src.shape[0], src.shape[1], src.strides[0], ch,
                                   _ptr(mapx, C.c_void_p), _ptr(mapy, C.c_void_p), dc,
                                   _ptr(dst, C.c_void_p), dr, dc, dst.strides[0])
        return dst

    def gray(self, src: np.ndarray, rgb: bool) -> np.ndarray:
        """cv2.cvtColor(src, COLOR_{RGB,BGR}[A]2GRAY)."""
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.empty(src.shape[:2], np.uint8)
        self.L.orc_cvt_gray_u8(_ptr(src, C.c_void_p), src.shape[0], src.shape[1], src.strides[0], src.shape[2],
                               int(bool(rgb)), _ptr(dst, C.c_void_p), dst.strides[0])
        return dst

    def undistort_points(self, xy: np.ndarray, fx, fy, cx, cy, dist) -> np.ndarray:
        """cv2.undistortPoints(xy, K, dist, None, K) for n x 2 float32 points; dist = (k1, k2, p1, p2[, k3])."""
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        d = np.zeros(5, np.float32); d[:len(dist)] = dist
        out = np.empty_like(xy)
        self.L.orc_undistort_points(xy.ctypes.data, len(xy), fx, fy, cx, cy, d.ctypes.data, out.ctypes.data)
        return out

    def prepare(self, src: np.ndarray, mapx=None, mapy=None, rgb: bool = False) -> np.ndarray:
        """The frame the extractor sees: remap (stereo_euroc.cc:168) then cvtColor (Tracking.cc:1264-1290)."""
        out = np.ascontiguousarray(src, np.uint8)
        if mapx is not None:
            out = self.remap(out, mapx, mapy)
        if out.ndim == 3:
            out = self.gray(out, rgb)
        return out

    def knn2(self, q: np.ndarray, t: np.ndarray):
        q = np.ascontiguousarray(q, np.uint8); t = np.ascontiguousarray(t, np.uint8)
        idx = np.empty((q.shape[0], 2), np.int32); dist = np.empty((q.shape[0], 2), np.int32)
        self.L.orc_bf_knn2(q.ctypes.data, q.shape[0], t.ctypes.data, t.shape[0], idx.ctypes.data, dist.ctypes.data)
        return idx, dist

    def hamming(self, a: np.ndarray, b: np.ndarray) -> int:
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.L.orc_hamming256(a.ctypes.data, b.ctypes.data)

    def distribute(self, cands: np.ndarray, minX, maxX, minY, maxY, N, tie_mode=0) -> np.ndarray:
        cands = np.ascontiguousarray(cands, CAND_DTYPE)
        out = np.empty(len(cands) + 8, CAND_DTYPE)
        n = self.L.orc_distribute_octree(cands.ctypes.data, len(cands), minX, maxX, minY, maxY, N, tie_mode,
                                         out.ctypes.data, len(out))
        if n < 0:
            raise ValueError("orc_distribute_octree failed")
        return out[:n].copy()


class Extractor:
    """Mirror of ORB_SLAM3::ORBextractor on the oracle (include/ORBextractor.h:49-83)."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 oracle: Oracle | None = None, tie_mode=0, trig_mode=0):
        self.o = oracle or Oracle()
        self.L = self.o.L
        self.h = self.L.orc_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        if not self.h:
            raise ValueError("orc_create failed")
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self.L.orc_set_modes(self.h, tie_mode, trig_mode)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def _farr(self, fn):
        return np.ctypeslib.as_array(fn(self.h), shape=(self.nlevels,)).copy()

    @property
    def scale_factors(self): return self._farr(self.L.orc_scale_factors)
    @property
    def inv_scale_factors(self): return self._farr(self.L.orc_inv_scale_factors)
    @property
    def level_sigma2(self): return self._farr(self.L.orc_level_sigma2)
    @property
    def inv_level_sigma2(self): return self._farr(self.L.orc_inv_level_sigma2)
    @property
    def features_per_level(self): return np.ctypeslib.as_array(self.L.orc_features_per_level(self.h), shape=(self.nlevels,)).copy()
    @property
    def umax(self): return np.ctypeslib.as_array(self.L.orc_umax(self.h), shape=(16,)).copy()

    def __call__(self, image: np.ndarray, lapping=(0, 0)):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32])."""
        if image is None or image.size == 0:
            return -1, np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1
        cap = self.nfeatures + 4 * self.nlevels + 64
        while True:
            kps = np.empty(cap, KP_DTYPE); desc = np.empty((cap, 32), np.uint8)
            n = C.c_int32(0)
            ret = self.L.orc_extract(self.h, image.ctypes.data, image.shape[0], image.shape[1], image.strides[0],
                                     int(lapping[0]), int(lapping[1]), kps.ctypes.data, desc.ctypes.data, cap, C.byref(n))
            if ret == -2:
                cap = n.value + 16
                continue
            if ret < 0:
                raise RuntimeError(f"orc_extract failed: {ret}")
            return ret, kps[:n.value].copy(), desc[:n.value].copy()

    def level_dims(self, l):
        r = C.c_int32(); c = C.c_int32()
        self.L.orc_level_dims(self.h, l, C.byref(r), C.byref(c))
        return r.value, c.value

    def level_plane(self, l) -> np.ndarray:
        """Padded plane (rows+38, cols+38) of mvImagePyramid[l]'s backing buffer."""
        st = C.c_size_t()
        p = self.L.orc_level_plane(self.h, l, C.byref(st))
        r, c = self.level_dims(l)
        buf = (C.c_uint8 * ((r + 38) * st.value)).from_address(p)
        return np.frombuffer(buf, np.uint8).reshape(r + 38, st.value)[:, :c + 38].copy()

    def level_blurred(self, l):
        st = C.c_size_t()
        p = self.L.orc_level_blurred(self.h, l, C.byref(st))
        if not p:
            return None
        r, c = self.level_dims(l)
        buf = (C.c_uint8 * (r * c)).from_address(p)
        return np.frombuffer(buf, np.uint8).reshape(r, c).copy()

    def level_candidates(self, l) -> np.ndarray:
        p = C.c_void_p()
        n = self.L.orc_level_candidates(self.h, l, C.byref(p))
        if n == 0:
            return np.empty(0, CAND_DTYPE)
        buf = (C.c_uint8 * (n * 12)).from_address(p.value)
        return np.frombuffer(buf, CAND_DTYPE).copy()

    def level_selected(self, l):
        p = C.c_void_p(); a = C.c_void_p()
        n = self.L.orc_level_selected(self.h, l, C.byref(p), C.byref(a))
        if n == 0:
            return np.empty(0, CAND_DTYPE), np.empty(0, np.float32)
        buf = (C.c_uint8 * (n * 12)).from_address(p.value)
        abuf = (C.c_float * n).from_address(a.value)
        return np.frombuffer(buf, CAND_DTYPE).copy(), np.frombuffer(abuf, np.float32).copy()


def stereo_match(eL: Extractor, eR: Extractor, kpL, descL, kpR, descR, mb: float, mbf: float):
    """Frame::ComputeStereoMatches on the pyramids left in eL/eR by their last __call__."""
    kpL = np.ascontiguousarray(kpL, KP_DTYPE); kpR = np.ascontiguousarray(kpR, KP_DTYPE)
    descL = np.ascontiguousarray(descL, np.uint8); descR = np.ascontiguousarray(descR, np.uint8)
    n = len(kpL)
    uR = np.empty(n, np.float32); depth = np.empty(n, np.float32)
    bd = np.empty(n, np.int32); bs = np.empty(n, np.int32)
    eL.L.orc_stereo_match(eL.h, eR.h, kpL.ctypes.data, n, descL.ctypes.data, kpR.ctypes.data, len(kpR),
                          descR.ctypes.data, mb, mbf, uR.ctypes.data, depth.ctypes.data, bd.ctypes.data, bs.ctypes.data)
    return uR, depth, bd, bs


class Grid:
    def __init__(self, o: Oracle, kps, minX, maxX, minY, maxY):
        self.L = o.L
        self.kps = np.ascontiguousarray(kps, KP_DTYPE)
        self.g = self.L.orc_grid_build(self.kps.ctypes.data, len(self.kps), minX, maxX, minY, maxY)

    def __del__(self):
        if getattr(self, "g", None):
            self.L.orc_grid_destroy(self.g)
            self.g = None

    def query(self, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.empty(len(self.kps) + 1, np.int32)
        n = self.L.orc_grid_query(self.g, self.kps.ctypes.data, x, y, r, minLevel, maxLevel, out.ctypes.data, len(out))
        return out[:n].copy()


def search_by_projection(grid: Grid, desc, uRight, scale_factors, mp: dict, th: float, nnratio: float,
                         far: bool = False, th_far: float = 0.0, assign=None):
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(grid.kps)
    assign = np.full(n, -1, np.int32) if assign is None else np.ascontiguousarray(assign, np.int32).copy()
    sf = np.ascontiguousarray(scale_factors, np.float32)
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, px, py, pxr = f("desc", np.uint8), f("projx", np.float32), f("projy", np.float32), f("projxr", np.float32)
    lv, vc, dp, iv = f("level", np.int32), f("viewcos", np.float32), f("depth", np.float32), f("inview", np.uint8)
    nm = grid.L.orc_search_by_projection(grid.g, grid.kps.ctypes.data, desc.ctypes.data, n,
                                         None if ur is None else ur.ctypes.data, sf.ctypes.data,
                                         len(px), md.ctypes.data, px.ctypes.data, py.ctypes.data, pxr.ctypes.data,
                                         lv.ctypes.data, vc.ctypes.data, dp.ctypes.data, iv.ctypes.data,
                                         th, nnratio, int(far), th_far, assign.ctypes.data)
    return nm, assign


def _p(a):
    return None if a is None else a.ctypes.data


def three_maxima(o: Oracle, sizes):
    sizes = np.ascontiguousarray(sizes, np.int32)
    out = np.zeros(3, np.int32)
    o.L.orc_three_maxima(sizes.ctypes.data, len(sizes), out[0:].ctypes.data, out[1:].ctypes.data, out[2:].ctypes.data)
    return tuple(int(v) for v in out)


def search_by_projection_frame(grid: Grid, desc, uRight, scale_factors, last: dict, th: float, mbf: float,
                               forward: bool, backward: bool, check_orientation: bool = True, cur_held=None):
    """last: dict of SoA arrays desc,u,v,invz,octave,angle,valid,obs (see orc_match.c)."""
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(grid.kps)
    assign = np.full(n, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    held = None if cur_held is None else np.ascontiguousarray(cur_held, np.uint8)
    f = lambda k, t: np.ascontiguousarray(last[k], t)
    ld, lu, lv, lz = f("desc", np.uint8), f("u", np.float32), f("v", np.float32), f("invz", np.float32)
    lo, la, lval, lobs = f("octave", np.int32), f("angle", np.float32), f("valid", np.uint8), f("obs", np.uint8)
    nm = grid.L.orc_search_by_projection_frame(grid.g, grid.kps.ctypes.data, desc.ctypes.data, n, _p(ur), sf.ctypes.data,
                                               len(lu), ld.ctypes.data, lu.ctypes.data, lv.ctypes.data, lz.ctypes.data,
                                               lo.ctypes.data, la.ctypes.data, lval.ctypes.data, lobs.ctypes.data,
                                               th, mbf, int(forward), int(backward), int(check_orientation),
                                               _p(held), assign.ctypes.data)
    return nm, assign


def search_for_initialization(o: Oracle, kps1, desc1, grid2: Grid, desc2, prev_matched, window: int = 100,
                              nnratio: float = 0.9, check_orientation: bool = True):
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE)
    desc1 = np.ascontiguousarray(desc1, np.uint8)
    desc2 = np.ascontiguousarray(desc2, np.uint8)
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.full(len(kps1), -1, np.int32)
    nm = o.L.orc_search_for_initialization(kps1.ctypes.data, desc1.ctypes.data, len(kps1), grid2.g,
                                           grid2.kps.ctypes.data, desc2.ctypes.data, len(grid2.kps),
                                           pm.ctypes.data, window, nnratio, int(check_orientation), m12.ctypes.data)
    return nm, m12, pm


def search_for_triangulation(o: Oracle, kf1: dict, kf2: dict, sf2, sigma2_2, ep, F12, only_stereo=False, coarse=False,
                             check_orientation=True, variant=0):
    """kfN: dict kps, desc, has_mp, uRight (or None), fv_node, fv_start, fv_idx."""
    def prep(k):
        return (np.ascontiguousarray(k["kps"], KP_DTYPE), np.ascontiguousarray(k["desc"], np.uint8),
                np.ascontiguousarray(k["has_mp"], np.uint8),
                None if k.get("uRight") is None else np.ascontiguousarray(k["uRight"], np.float32),
                np.ascontiguousarray(k["fv_node"], np.int32), np.ascontiguousarray(k["fv_start"], np.int32),
                np.ascontiguousarray(k["fv_idx"], np.int32))
    a, b = prep(kf1), prep(kf2)
    sf2 = np.ascontiguousarray(sf2, np.float32); sg = np.ascontiguousarray(sigma2_2, np.float32)
    ep = np.ascontiguousarray(ep, np.float32); F = np.ascontiguousarray(F12, np.float32).reshape(9)
    m12 = np.full(len(a[0]), -1, np.int32)
    args = []
    for k in (a, b):
        args += [k[0].ctypes.data, k[1].ctypes.data, len(k[0]), k[2].ctypes.data, _p(k[3]),
                 k[4].ctypes.data, k[5].ctypes.data, k[6].ctypes.data, len(k[4])]
    nm = o.L.orc_search_for_triangulation(*args, sf2.ctypes.data, sg.ctypes.data, ep.ctypes.data, F.ctypes.data,
                                          int(only_stereo), int(coarse), int(check_orientation), int(variant),
                                          m12.ctypes.data)
    return nm, m12


def search_by_bow(o: Oracle, mode: int, kf1: dict, kf2: dict, nnratio: float, check_orientation: bool = True):
    """kfN: dict kps, desc, valid (or None for side 2 in mode 0), fv_node, fv_start, fv_idx."""
    def prep(k):
        return (np.ascontiguousarray(k["kps"], KP_DTYPE), np.ascontiguousarray(k["desc"], np.uint8),
                None if k.get("valid") is None else np.ascontiguousarray(k["valid"], np.uint8),
                np.ascontiguousarray(k["fv_node"], np.int32), np.ascontiguousarray(k["fv_start"], np.int32),
                np.ascontiguousarray(k["fv_idx"], np.int32))
    a, b = prep(kf1), prep(kf2)
    out = np.full(len(b[0]) if mode == 0 else len(a[0]), -1, np.int32)
    args = []
    for k in (a, b):
        args += [k[0].ctypes.data, k[1].ctypes.data, len(k[0]), _p(k[2]), k[3].ctypes.data, k[4].ctypes.data,
                 k[5].ctypes.data, len(k[3])]
    nm = o.L.orc_search_by_bow(int(mode), *args, nnratio, int(check_orientation), out.ctypes.data)
    return nm, out


def search_by_projection_keyframe(grid: Grid, desc, scale_factors, mp: dict, th: float, orb_dist: int,
                                  check_orientation: bool = True, cur_held=None):
    """mp: SoA dict desc,u,v,level,angle,valid."""
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(grid.kps)
    assign = np.full(n, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    held = None if cur_held is None else np.ascontiguousarray(cur_held, np.uint8)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, mu, mv = f("desc", np.uint8), f("u", np.float32), f("v", np.float32)
    ml, ma, mval = f("level", np.int32), f("angle", np.float32), f("valid", np.uint8)
    nm = grid.L.orc_search_by_projection_keyframe(grid.g, grid.kps.ctypes.data, desc.ctypes.data, n, sf.ctypes.data,
                                                  len(mu), md.ctypes.data, mu.ctypes.data, mv.ctypes.data, ml.ctypes.data,
                                                  ma.ctypes.data, mval.ctypes.data, th, int(orb_dist), int(check_orientation),
                                                  _p(held), assign.ctypes.data)
    return nm, assign


def search_by_projection_sim3(grid: Grid, desc, scale_factors, mp: dict, th: float, ratio_hamming: float, matched=None):
    """ORBmatcher::SearchByProjection(KeyFrame*, Scw, ...) (src/ORBmatcher.cc:549, :681).  mp: SoA dict desc,u,v,level,valid."""
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(grid.kps)
    assign = np.full(n, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    held = None if matched is None else np.ascontiguousarray(matched, np.uint8)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, mu, mv, ml, mval = f("desc", np.uint8), f("u", np.float32), f("v", np.float32), f("level", np.int32), f("valid", np.uint8)
    nm = grid.L.orc_search_by_projection_sim3(grid.g, grid.kps.ctypes.data, desc.ctypes.data, n, sf.ctypes.data, len(mu),
                                              md.ctypes.data, mu.ctypes.data, mv.ctypes.data, ml.ctypes.data, mval.ctypes.data,
                                              th, ratio_hamming, _p(held), assign.ctypes.data)
    return nm, assign


def fuse(grid: Grid, desc, uRight, scale_factors, inv_sigma2, mp: dict, th: float):
    """The search of ORBmatcher::Fuse (src/ORBmatcher.cc:1823, :2051).  mp: desc,u,v,level,valid and optionally ur."""
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(grid.kps)
    sf = np.ascontiguousarray(scale_factors, np.float32); isg = np.ascontiguousarray(inv_sigma2, np.float32)
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, mu, mv, ml, mval = f("desc", np.uint8), f("u", np.float32), f("v", np.float32), f("level", np.int32), f("valid", np.uint8)
    mur = f("ur", np.float32) if mp.get("ur") is not None else None
    bi = np.empty(len(mu), np.int32); bd = np.empty(len(mu), np.int32)
    nf = grid.L.orc_fuse(grid.g, grid.kps.ctypes.data, desc.ctypes.data, n, _p(ur), sf.ctypes.data, isg.ctypes.data, len(mu),
                         md.ctypes.data, mu.ctypes.data, mv.ctypes.data, _p(mur), ml.ctypes.data, mval.ctypes.data, th,
                         bi.ctypes.data, bd.ctypes.data)
    return nf, bi, bd


def search_by_sim3(grid1: Grid, desc1, grid2: Grid, desc2, scale_factors, rows1: dict, rows2: dict, th: float):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:2201).  rows1: KF1's map points projected into KF2 (desc,u,v,level,valid)."""
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    n1, n2 = len(grid1.kps), len(grid2.kps)
    g = lambda r, k, t: np.ascontiguousarray(r[k], t)
    a = [g(rows1, "desc", np.uint8), g(rows1, "u", np.float32), g(rows1, "v", np.float32), g(rows1, "level", np.int32), g(rows1, "valid", np.uint8)]
    b = [g(rows2, "desc", np.uint8), g(rows2, "u", np.float32), g(rows2, "v", np.float32), g(rows2, "level", np.int32), g(rows2, "valid", np.uint8)]
    m12 = np.empty(n1, np.int32)
    nf = grid1.L.orc_search_by_sim3(grid1.g, grid1.kps.ctypes.data, desc1.ctypes.data, n1, grid2.g, grid2.kps.ctypes.data, desc2.ctypes.data, n2,
                                    sf.ctypes.data, *[x.ctypes.data for x in a], *[x.ctypes.data for x in b], th, m12.ctypes.data)
    return nf, m12


def is_in_frustum(o: Oracle, pose: dict, cam: dict, bounds, log_sf: float, nlevels: int, pts: dict, cos_limit: float = 0.5):
    """Frame::isInFrustum (src/Frame.cc:603-711) for every point.  pose: Rcw(3x3), tcw(3), Ow(3); cam: fx, fy, cx, cy, mbf;
    pts: pos(n,3), normal(n,3), min_dist, max_dist, candidate.  Returns (nToMatch, dict of the mTrack* fields)."""
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    R, t, Ow = f32(pose["Rcw"]).reshape(9), f32(pose["tcw"]), f32(pose["Ow"])
    pos, nrm, dmin, dmax = f32(pts["pos"]), f32(pts["normal"]), f32(pts["min_dist"]), f32(pts["max_dist"])
    cand = np.ascontiguousarray(pts["candidate"], np.uint8)
    n = len(dmin)
    out = dict(inview=np.empty(n, np.uint8), projx=np.empty(n, np.float32), projy=np.empty(n, np.float32), projxr=np.empty(n, np.float32),
               depth=np.empty(n, np.float32), viewcos=np.empty(n, np.float32), level=np.empty(n, np.int32))
    nin = o.L.orc_is_in_frustum(R.ctypes.data, t.ctypes.data, Ow.ctypes.data, cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["mbf"],
                                bounds[0], bounds[1], bounds[2], bounds[3], log_sf, nlevels, n, pos.ctypes.data, nrm.ctypes.data,
                                dmin.ctypes.data, dmax.ctypes.data, cand.ctypes.data, cos_limit,
                                *[out[k].ctypes.data for k in ("inview", "projx", "projy", "projxr", "depth", "viewcos", "level")])
    return nin, out


def distinctive_descriptors(o: Oracle, start, desc):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:450-531) per CSR group.  Returns (best_idx, best_median)."""
    start = np.ascontiguousarray(start, np.int32); desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(start) - 1
    bi = np.empty(n, np.int32); bm = np.empty(n, np.int32)
    o.L.orc_distinctive_descriptors(n, start.ctypes.data, desc.ctypes.data, bi.ctypes.data, bm.ctypes.data)
    return bi, bm


def vocabulary_csr(parent, is_leaf):
    """Children lists / word ids as TemplatedVocabulary::loadFromTextFile builds them (TemplatedVocabulary.h:1338-1415):
    node i+1 is row i of the file; children in file order; word ids number the leaves in file order."""
    parent = np.asarray(parent, np.int64); is_leaf = np.asarray(is_leaf, bool)
    nn = len(parent) + 1
    ids = np.arange(1, nn)
    order = np.argsort(parent, kind="stable")
    child_idx = ids[order].astype(np.int32)
    counts = np.bincount(parent, minlength=nn)
    child_start = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    word = np.zeros(nn, np.int32)
    word[1:][is_leaf] = np.arange(int(is_leaf.sum()), dtype=np.int32)
    return child_start, child_idx, word


def bow_transform(o: Oracle, voc: dict, desc, levelsup: int = 4):
    """voc: parent, is_leaf, desc (nn-1 x 32), weight (nn-1), L.  Returns per feature (word, weight, node)."""
    cs, ci, word = vocabulary_csr(voc["parent"], voc["is_leaf"])
    nd = np.concatenate([np.zeros((1, 32), np.uint8), np.ascontiguousarray(voc["desc"], np.uint8)])
    nw = np.concatenate([[0.0], np.asarray(voc["weight"], np.float64)])
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(desc)
    w = np.empty(n, np.int32); wt = np.empty(n, np.float64); nid = np.empty(n, np.int32)
    o.L.orc_bow_transform(cs.ctypes.data, ci.ctypes.data, nd.ctypes.data, nw.ctypes.data, word.ctypes.data, int(voc["L"]), levelsup,
                          n, desc.ctypes.data, w.ctypes.data, wt.ctypes.data, nid.ctypes.data)
    return w, wt, nid


def compute_bow(o: Oracle, voc: dict, desc, levelsup: int = 4):
    """Frame::ComputeBoW (src/Frame.cc:905-922): (bow_id, bow_val, fv_node, fv_start, fv_idx)."""
    w, wt, nid = bow_transform(o, voc, desc, levelsup)
    n = len(w)
    bid = np.empty(n, np.int32); bval = np.empty(n, np.float64)
    fn = np.empty(n, np.int32); fs = np.empty(n + 1, np.int32); fi = np.empty(n, np.int32)
    nfv = C.c_int(0)
    nw = o.L.orc_bow_vectors(n, w.ctypes.data, wt.ctypes.data, nid.ctypes.data, bid.ctypes.data, bval.ctypes.data,
                             fn.ctypes.data, fs.ctypes.data, fi.ctypes.data, C.byref(nfv))
    k = nfv.value
    return bid[:nw].copy(), bval[:nw].copy(), fn[:k].copy(), fs[:k + 1].copy(), fi[:fs[k]].copy()


def search_by_projection_rig(gridL: Grid, gridR: Grid, desc, l2r, r2l, scale_factors, mp: dict, th: float, nnratio: float,
                             far: bool = False, th_far: float = 0.0, assign=None):
    """SearchByProjection(Frame&, vector<MapPoint*>&, ...) for Nleft != -1 (src/ORBmatcher.cc:45-239).
    mp: desc, depth, and per side inview / projx / projy / level / viewcos with suffix L or R."""
    desc = np.ascontiguousarray(desc, np.uint8)
    nL, nR = len(gridL.kps), len(gridR.kps)
    assign = np.full(nL + nR, -1, np.int32) if assign is None else np.ascontiguousarray(assign, np.int32).copy()
    sf = np.ascontiguousarray(scale_factors, np.float32)
    l2r = np.ascontiguousarray(l2r, np.int32); r2l = np.ascontiguousarray(r2l, np.int32)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    args = [f("desc", np.uint8), f("depth", np.float32)]
    for s in ("L", "R"):
        args += [f("inview" + s, np.uint8), f("projx" + s, np.float32), f("projy" + s, np.float32), f("level" + s, np.int32), f("viewcos" + s, np.float32)]
    nm = gridL.L.orc_search_by_projection_rig(gridL.g, gridL.kps.ctypes.data, nL, gridR.g, gridR.kps.ctypes.data, nR, desc.ctypes.data,
                                              l2r.ctypes.data, r2l.ctypes.data, sf.ctypes.data, len(args[1]), *[a.ctypes.data for a in args],
                                              th, nnratio, int(far), th_far, assign.ctypes.data)
    return nm, assign


# ---- two-camera rig branches, callback-driven triangulation searches, fisheye stereo (orc_match.c, bottom) ----
VERDICT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int)


def _kf_args(k, with_valid_key):
    kps = np.ascontiguousarray(k["kps"], KP_DTYPE); desc = np.ascontiguousarray(k["desc"], np.uint8)
    flag = None if k.get(with_valid_key) is None else np.ascontiguousarray(k[with_valid_key], np.uint8)
    fn, fs, fi = (np.ascontiguousarray(k[x], np.int32) for x in ("fv_node", "fv_start", "fv_idx"))
    return kps, desc, flag, fn, fs, fi


def search_by_projection_frame_rig(gridL: Grid, gridR: Grid, desc, scale_factors, last: dict, th: float,
                                   forward: bool, backward: bool, check_orientation: bool = True, cur_held=None):
    """last: SoA dict desc,u,v,ur,vr,invz,octave,angle,valid,obs.  Returns (nmatches, assign[nL + nR])."""
    L = gridL.L
    L.orc_search_by_projection_frame_rig.restype = C.c_int
    L.orc_search_by_projection_frame_rig.argtypes = [C.c_void_p, C.c_void_p, C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_int] + \
        [C.c_void_p] * 10 + [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    desc = np.ascontiguousarray(desc, np.uint8)
    nL, nR = len(gridL.kps), len(gridR.kps)
    assign = np.full(nL + nR, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    held = None if cur_held is None else np.ascontiguousarray(cur_held, np.uint8)
    f = lambda k, t: np.ascontiguousarray(last[k], t)
    ld = f("desc", np.uint8)
    lu, lv, lur, lvr, lz = (f(k, np.float32) for k in ("u", "v", "ur", "vr", "invz"))
    lo, la, lval, lobs = f("octave", np.int32), f("angle", np.float32), f("valid", np.uint8), f("obs", np.uint8)
    nm = L.orc_search_by_projection_frame_rig(gridL.g, gridL.kps.ctypes.data, nL, gridR.g, gridR.kps.ctypes.data, nR,
                                              desc.ctypes.data, sf.ctypes.data, len(lu), ld.ctypes.data, lu.ctypes.data, lv.ctypes.data,
                                              lur.ctypes.data, lvr.ctypes.data, lz.ctypes.data, lo.ctypes.data, la.ctypes.data,
                                              lval.ctypes.data, lobs.ctypes.data, th, int(forward), int(backward),
                                              int(check_orientation), _p(held), assign.ctypes.data)
    return nm, assign


def search_by_bow_rig(o: Oracle, kf1: dict, kf2: dict, nleft2: int, nnratio: float, check_orientation: bool = True):
    L = o.L
    L.orc_search_by_bow_rig.restype = C.c_int
    L.orc_search_by_bow_rig.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_float, C.c_int, C.c_void_p]
    a, b = _kf_args(kf1, "valid"), _kf_args(kf2, "valid")
    out = np.full(len(b[0]), -1, np.int32)
    nm = L.orc_search_by_bow_rig(a[0].ctypes.data, a[1].ctypes.data, len(a[0]), a[2].ctypes.data, a[3].ctypes.data, a[4].ctypes.data,
                                 a[5].ctypes.data, len(a[3]), b[0].ctypes.data, b[1].ctypes.data, len(b[0]), int(nleft2),
                                 b[3].ctypes.data, b[4].ctypes.data, b[5].ctypes.data, len(b[3]), nnratio, int(check_orientation),
                                 out.ctypes.data)
    return nm, out


def search_for_triangulation_cb(o: Oracle, overload: int, kf1: dict, kf2: dict, verdict, only_stereo=False, coarse=False,
                                check_orientation=True):
    """verdict(idx1, idx2) -> bool, a Python callable.  kfN: dict kps, desc, has_mp, fv_node, fv_start, fv_idx."""
    L = o.L
    L.orc_search_for_triangulation_cb.restype = C.c_int
    L.orc_search_for_triangulation_cb.argtypes = [C.c_int] + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int] * 2 + \
        [C.c_int, C.c_int, C.c_int, VERDICT_FN, C.c_void_p, C.c_void_p]
    a, b = _kf_args(kf1, "has_mp"), _kf_args(kf2, "has_mp")
    m12 = np.full(len(a[0]), -1, np.int32)
    cb = VERDICT_FN(lambda user, i1, i2: int(bool(verdict(i1, i2))))
    args = []
    for k in (a, b):
        args += [k[0].ctypes.data, k[1].ctypes.data, len(k[0]), k[2].ctypes.data, k[3].ctypes.data, k[4].ctypes.data, k[5].ctypes.data, len(k[3])]
    nm = L.orc_search_for_triangulation_cb(int(overload), *args, int(only_stereo), int(coarse), int(check_orientation), cb, None, m12.ctypes.data)
    return nm, m12


def stereo_fisheye_match(o: Oracle, descL, monoLeft: int, descR, monoRight: int):
    """Frame::ComputeStereoFishEyeMatches up to the triangulation gate.  Returns (good, l2r, r2l, best_dist)."""
    L = o.L
    L.orc_stereo_fisheye_match.restype = C.c_int
    L.orc_stereo_fisheye_match.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    descL = np.ascontiguousarray(descL, np.uint8); descR = np.ascontiguousarray(descR, np.uint8)
    l2r = np.full(len(descL), -1, np.int32); r2l = np.full(len(descR), -1, np.int32); bd = np.full(len(descL), -1, np.int32)
    good = L.orc_stereo_fisheye_match(descL.ctypes.data, len(descL), int(monoLeft), descR.ctypes.data, len(descR), int(monoRight),
                                      l2r.ctypes.data, r2l.ctypes.data, bd.ctypes.data)
    return good, l2r, r2l, bd

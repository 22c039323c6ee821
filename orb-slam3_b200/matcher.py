"""Python mirrors of Frame::ComputeStereoMatches, BFMatcher kNN and ORBmatcher::SearchByProjection
on the C ABI (include/orbx.h).  All compute happens in liborbx.so on the GPU."""
from __future__ import annotations

import numpy as np

from . import binding
from .binding import KP_DTYPE, check


def ComputeStereoMatches(extL, extR, mb: float, mbf: float, frame: int | None = 0):
    """Frame::ComputeStereoMatches (src/Frame.cc:1027-1276) on the device-resident results of the
    preceding extraction on extL / extR.  frame=None -> all frames: returns ([B,cap], [B,cap])."""
    lib = binding.load()
    cap = extL.max_keypoints(*extL_shape(extL))
    if frame is None:
        B = extL_frames(extL)
        uR = np.empty((B, cap), np.float32); depth = np.empty((B, cap), np.float32)
        check(lib.orbx_stereo_match_batch(extL.h, extR.h, mb, mbf, uR.ctypes.data, depth.ctypes.data, cap))
        return uR, depth
    uR = np.empty(cap, np.float32); depth = np.empty(cap, np.float32)
    check(lib.orbx_stereo_match(extL.h, extR.h, frame, mb, mbf, uR.ctypes.data, depth.ctypes.data, cap))
    return uR, depth


def extL_shape(ext):
    return ext._last_shape[1:]


def extL_frames(ext):
    return ext._last_shape[0]


def knn2_hamming(q: np.ndarray, t: np.ndarray, device: int = 0):
    """cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, 2) (src/Frame.cc:1458)."""
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32); t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    idx = np.empty((len(q), 2), np.int32); dist = np.empty((len(q), 2), np.int32)
    check(binding.load().orbx_knn2_hamming(device, q.ctypes.data, len(q), t.ctypes.data, len(t), idx.ctypes.data, dist.ctypes.data))
    return idx, dist


def SearchByProjection(kps, desc, uRight, bounds, scale_factors, mp: dict, th: float, nnratio: float,
                       far: bool = False, th_far: float = 0.0, assign=None, device: int = 0):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFar, thFar) (src/ORBmatcher.cc:45).
    Returns (nmatches, assign)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps)
    assign = np.full(n, -1, np.int32) if assign is None else np.ascontiguousarray(assign, np.int32).copy()
    sf = np.ascontiguousarray(scale_factors, np.float32)
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, px, py, pxr = f("desc", np.uint8), f("projx", np.float32), f("projy", np.float32), f("projxr", np.float32)
    lv, vc, dp, iv = f("level", np.int32), f("viewcos", np.float32), f("depth", np.float32), f("inview", np.uint8)
    nm = check(binding.load().orbx_search_by_projection(
        device, kps.ctypes.data, desc.ctypes.data, n, None if ur is None else ur.ctypes.data,
        bounds[0], bounds[1], bounds[2], bounds[3], sf.ctypes.data, len(sf),
        len(px), md.ctypes.data, px.ctypes.data, py.ctypes.data, pxr.ctypes.data, lv.ctypes.data, vc.ctypes.data,
        dp.ctypes.data, iv.ctypes.data, th, nnratio, int(far), th_far, assign.ctypes.data))
    return nm, assign

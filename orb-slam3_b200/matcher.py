"""Python mirrors of Frame::ComputeStereoMatches, BFMatcher kNN and the ORBmatcher searches
(SearchByProjection x2, SearchForInitialization, SearchForTriangulation[_], ComputeThreeMaxima)
on the C ABI (include/orbx.h).  All compute happens in liborbx.so on the GPU."""
from __future__ import annotations

import ctypes as C

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


def ExtractStereo(extL, extR, imL: np.ndarray, imR: np.ndarray, mb: float, mbf: float):
    """The stereo Frame constructor's hot part (src/Frame.cc:132-137 two ExtractORB threads + :154 ComputeStereoMatches)
    as one call: returns (kpsL, descL, kpsR, descR, mvuRight, mvDepth)."""
    import ctypes as C
    for im in (imL, imR):
        assert im.dtype == np.uint8 and im.ndim == 2 and im.strides[1] == 1, "CV_8UC1 expected"
    assert imL.shape == imR.shape
    rows, cols = imL.shape
    cap = extL.max_keypoints(rows, cols)
    kL = np.empty(cap, KP_DTYPE); dL = np.empty((cap, 32), np.uint8)
    kR = np.empty(cap, KP_DTYPE); dR = np.empty((cap, 32), np.uint8)
    uR = np.empty(cap, np.float32); depth = np.empty(cap, np.float32)
    nL = C.c_int(0); nR = C.c_int(0)
    check(binding.load().orbx_extract_stereo(extL.h, extR.h, imL.ctypes.data, imL.strides[0], imR.ctypes.data, imR.strides[0],
                                             rows, cols, kL.ctypes.data, dL.ctypes.data, C.byref(nL),
                                             kR.ctypes.data, dR.ctypes.data, C.byref(nR), cap, mb, mbf,
                                             uR.ctypes.data, depth.ctypes.data))
    extL._last_shape = extR._last_shape = (1, rows, cols)
    a, b = nL.value, nR.value
    return kL[:a].copy(), dL[:a].copy(), kR[:b].copy(), dR[:b].copy(), uR[:a].copy(), depth[:a].copy()


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


def _p(a):
    return None if a is None else a.ctypes.data


def GetFeaturesInArea(kps, desc, bounds, q_desc, q_x, q_y, q_r, q_min, q_max, cap: int | None = None, device: int = 0):
    """Frame::GetFeaturesInArea (src/Frame.cc:784-871) for many windows + DescriptorDistance of every hit.
    Returns (start[nq+1], idx, dist)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    qd = np.ascontiguousarray(q_desc, np.uint8)
    qx, qy, qr = (np.ascontiguousarray(a, np.float32) for a in (q_x, q_y, q_r))
    qmin, qmax = (np.ascontiguousarray(a, np.int32) for a in (q_min, q_max))
    nq = len(qx)
    cap = max(1, nq * max(1, len(kps))) if cap is None else cap
    start = np.zeros(nq + 1, np.int32); idx = np.empty(cap, np.int32); dist = np.empty(cap, np.int32)
    tot = check(binding.load().orbx_window_candidates(
        device, kps.ctypes.data, desc.ctypes.data, len(kps), bounds[0], bounds[1], bounds[2], bounds[3],
        nq, qd.ctypes.data, qx.ctypes.data, qy.ctypes.data, qr.ctypes.data, qmin.ctypes.data, qmax.ctypes.data,
        start.ctypes.data, idx.ctypes.data, dist.ctypes.data, cap))
    return start, idx[:tot].copy(), dist[:tot].copy()


def SearchByProjectionFrame(kps, desc, uRight, bounds, scale_factors, last: dict, th: float, mbf: float,
                            forward: bool, backward: bool, check_orientation: bool = True, cur_held=None, device: int = 0):
    """ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (src/ORBmatcher.cc:2469).
    last: SoA dict desc,u,v,invz,octave,angle,valid,obs.  Returns (nmatches, assign)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps)
    assign = np.full(n, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    held = None if cur_held is None else np.ascontiguousarray(cur_held, np.uint8)
    f = lambda k, t: np.ascontiguousarray(last[k], t)
    ld, lu, lv, lz = f("desc", np.uint8), f("u", np.float32), f("v", np.float32), f("invz", np.float32)
    lo, la, lval, lobs = f("octave", np.int32), f("angle", np.float32), f("valid", np.uint8), f("obs", np.uint8)
    nm = check(binding.load().orbx_search_by_projection_frame(
        device, kps.ctypes.data, desc.ctypes.data, n, _p(ur), bounds[0], bounds[1], bounds[2], bounds[3],
        sf.ctypes.data, len(sf), len(lu), ld.ctypes.data, lu.ctypes.data, lv.ctypes.data, lz.ctypes.data,
        lo.ctypes.data, la.ctypes.data, lval.ctypes.data, lobs.ctypes.data,
        th, mbf, int(forward), int(backward), int(check_orientation), _p(held), assign.ctypes.data))
    return nm, assign


def SearchByProjectionKeyFrame(kps, desc, bounds, scale_factors, mp: dict, th: float, orb_dist: int,
                               check_orientation: bool = True, cur_held=None, device: int = 0):
    """ORBmatcher::SearchByProjection(Frame& Current, KeyFrame*, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2723).
    mp: SoA dict desc,u,v,level,angle,valid.  Returns (nmatches, assign)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps)
    assign = np.full(n, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    held = None if cur_held is None else np.ascontiguousarray(cur_held, np.uint8)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, mu, mv = f("desc", np.uint8), f("u", np.float32), f("v", np.float32)
    ml, ma, mval = f("level", np.int32), f("angle", np.float32), f("valid", np.uint8)
    nm = check(binding.load().orbx_search_by_projection_keyframe(
        device, kps.ctypes.data, desc.ctypes.data, n, bounds[0], bounds[1], bounds[2], bounds[3], sf.ctypes.data, len(sf),
        len(mu), md.ctypes.data, mu.ctypes.data, mv.ctypes.data, ml.ctypes.data, ma.ctypes.data, mval.ctypes.data,
        th, int(orb_dist), int(check_orientation), _p(held), assign.ctypes.data))
    return nm, assign


def SearchForInitialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window: int = 100, nnratio: float = 0.9,
                            check_orientation: bool = True, device: int = 0):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:799).  Returns (nmatches, matches12, prev_matched)."""
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); desc1 = np.ascontiguousarray(desc1, np.uint8)
    kps2 = np.ascontiguousarray(kps2, KP_DTYPE); desc2 = np.ascontiguousarray(desc2, np.uint8)
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.full(len(kps1), -1, np.int32)
    nm = check(binding.load().orbx_search_for_initialization(
        device, kps1.ctypes.data, desc1.ctypes.data, len(kps1), kps2.ctypes.data, desc2.ctypes.data, len(kps2),
        bounds[0], bounds[1], bounds[2], bounds[3], pm.ctypes.data, window, nnratio, int(check_orientation), m12.ctypes.data))
    return nm, m12, pm


def SearchForTriangulation(kf1: dict, kf2: dict, sf2, sigma2_2, ep, F12, only_stereo: bool = False, coarse: bool = False,
                           check_orientation: bool = True, variant: int = 0, device: int = 0):
    """ORBmatcher::SearchForTriangulation (variant 0, src/ORBmatcher.cc:1107) / SearchForTriangulation_ (variant 1,
    :1388).  kfN: dict kps, desc, has_mp, uRight (or None), fv_node, fv_start, fv_idx.  Returns (nmatches, matches12)."""
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
    nm = check(binding.load().orbx_search_for_triangulation(
        device, *args, sf2.ctypes.data, sg.ctypes.data, len(sf2), ep.ctypes.data, F.ctypes.data,
        int(only_stereo), int(coarse), int(check_orientation), int(variant), m12.ctypes.data))
    return nm, m12


def SearchByBoW(mode: int, kf1: dict, kf2: dict, nnratio: float, check_orientation: bool = True, device: int = 0):
    """ORBmatcher::SearchByBoW — mode 0: (KeyFrame, Frame) src/ORBmatcher.cc:314, mode 1: (KeyFrame, KeyFrame) :955.
    kfN: dict kps, desc, valid (None allowed for the Frame side of mode 0), fv_node, fv_start, fv_idx.
    Returns (nmatches, matches) — mode 0: per Frame keypoint the KF keypoint index, mode 1: per KF1 keypoint the KF2 one."""
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
    nm = check(binding.load().orbx_search_by_bow(device, int(mode), *args, nnratio, int(check_orientation), out.ctypes.data))
    return nm, out


class ORBVocabulary:
    """ORBVocabulary (include/ORBVocabulary.h) on the device, built from the rows of its text file
    (TemplatedVocabulary::loadFromTextFile, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1415)."""

    def __init__(self, k: int, L: int, parent, is_leaf, desc, weight, device: int = 0):
        import ctypes as C
        self.lib = binding.load()
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); weight = np.ascontiguousarray(weight, np.float64)
        h = C.c_void_p()
        check(self.lib.orbx_vocabulary_create(device, k, L, len(parent), parent.ctypes.data, is_leaf.ctypes.data,
                                              desc.ctypes.data, weight.ctypes.data, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.orbx_vocabulary_destroy(self.h); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform(self, desc, levelsup: int = 4):
        """Per feature (word id, weight, node id) — TemplatedVocabulary::transform, :1218-1258."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        w = np.empty(n, np.int32); wt = np.empty(n, np.float64); nid = np.empty(n, np.int32)
        check(self.lib.orbx_bow_transform(self.h, desc.ctypes.data, n, levelsup, w.ctypes.data, wt.ctypes.data, nid.ctypes.data))
        return w, wt, nid

    def ComputeBoW(self, desc, levelsup: int = 4):
        """Frame::ComputeBoW (src/Frame.cc:905-922): (bow_id, bow_val, fv_node, fv_start, fv_idx)."""
        import ctypes as C
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        bid = np.empty(n, np.int32); bval = np.empty(n, np.float64)
        fn = np.empty(n, np.int32); fs = np.empty(n + 1, np.int32); fi = np.empty(n, np.int32)
        nb = C.c_int(0); nf = C.c_int(0)
        check(self.lib.orbx_compute_bow(self.h, desc.ctypes.data, n, levelsup, bid.ctypes.data, bval.ctypes.data, C.byref(nb),
                                        fn.ctypes.data, fs.ctypes.data, fi.ctypes.data, C.byref(nf)))
        k = nf.value
        return bid[:nb.value].copy(), bval[:nb.value].copy(), fn[:k].copy(), fs[:k + 1].copy(), fi[:fs[k]].copy()


def ComputeDistinctiveDescriptors(start, desc, device: int = 0):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:450-531), batched: CSR groups of descriptors.
    Returns (best_idx, best_median)."""
    start = np.ascontiguousarray(start, np.int32); desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(start) - 1
    bi = np.empty(n, np.int32); bm = np.empty(n, np.int32)
    check(binding.load().orbx_distinctive_descriptors(device, n, start.ctypes.data, desc.ctypes.data, bi.ctypes.data, bm.ctypes.data))
    return bi, bm


def SearchByProjectionRig(kpsL, kpsR, desc, l2r, r2l, bounds, scale_factors, mp: dict, th: float, nnratio: float,
                          far: bool = False, th_far: float = 0.0, assign=None, device: int = 0):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) for Nleft != -1 (src/ORBmatcher.cc:45-239).
    mp: desc, depth and per side inview / projx / projy / level / viewcos with suffix L or R.  Returns (nmatches, assign)."""
    kpsL = np.ascontiguousarray(kpsL, KP_DTYPE); kpsR = np.ascontiguousarray(kpsR, KP_DTYPE)
    desc = np.ascontiguousarray(desc, np.uint8)
    nL, nR = len(kpsL), len(kpsR)
    assign = np.full(nL + nR, -1, np.int32) if assign is None else np.ascontiguousarray(assign, np.int32).copy()
    sf = np.ascontiguousarray(scale_factors, np.float32)
    l2r = np.ascontiguousarray(l2r, np.int32); r2l = np.ascontiguousarray(r2l, np.int32)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    args = [f("desc", np.uint8), f("depth", np.float32)]
    for s in ("L", "R"):
        args += [f("inview" + s, np.uint8), f("projx" + s, np.float32), f("projy" + s, np.float32), f("level" + s, np.int32), f("viewcos" + s, np.float32)]
    nm = check(binding.load().orbx_search_by_projection_rig(
        device, kpsL.ctypes.data, nL, kpsR.ctypes.data, nR, desc.ctypes.data, l2r.ctypes.data, r2l.ctypes.data,
        bounds[0], bounds[1], bounds[2], bounds[3], sf.ctypes.data, len(sf), len(args[1]), *[a.ctypes.data for a in args],
        th, nnratio, int(far), th_far, assign.ctypes.data))
    return nm, assign


def UndistortKeyPoints(kps, fx, fy, cx, cy, dist_coef, device: int = 0):
    """Frame::UndistortKeyPoints (src/Frame.cc:924-960): mvKeysUn from mvKeys; dist_coef = (k1, k2, p1, p2[, k3])."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    d = np.ascontiguousarray(dist_coef, np.float32)
    out = np.empty_like(kps)
    check(binding.load().orbx_undistort_keypoints(device, kps.ctypes.data, len(kps), fx, fy, cx, cy, d.ctypes.data, len(d), out.ctypes.data))
    return out


def SearchLocalPoints(kps, desc, uRight, bounds, scale_factors, pose: dict, cam: dict, log_sf: float, pts: dict,
                      cos_limit: float = 0.5, th: float = 1.0, nnratio: float = 0.8, far: bool = False, th_far: float = 0.0,
                      assign=None, device: int = 0):
    """Tracking::SearchLocalPoints (src/Tracking.cc:3808-3893): Frame::isInFrustum on the device + SearchByProjection.
    pose: Rcw, tcw, Ow; cam: fx, fy, cx, cy, mbf; pts: pos(n,3), normal(n,3), min_dist, max_dist (mfMinDistance / mfMaxDistance as stored), desc, candidate.
    Returns (nmatches, assign, dict of the mTrack* fields)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps); sf = np.ascontiguousarray(scale_factors, np.float32)
    assign = np.full(n, -1, np.int32) if assign is None else np.ascontiguousarray(assign, np.int32).copy()
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    R, t, Ow = f32(pose["Rcw"]).reshape(9), f32(pose["tcw"]), f32(pose["Ow"])
    pos, nrm, dmin, dmax = f32(pts["pos"]), f32(pts["normal"]), f32(pts["min_dist"]), f32(pts["max_dist"])
    md = np.ascontiguousarray(pts["desc"], np.uint8)
    cand = None if pts.get("candidate") is None else np.ascontiguousarray(pts["candidate"], np.uint8)
    M = len(dmin)
    out = dict(inview=np.empty(M, np.uint8), projx=np.empty(M, np.float32), projy=np.empty(M, np.float32), projxr=np.empty(M, np.float32),
               depth=np.empty(M, np.float32), viewcos=np.empty(M, np.float32), level=np.empty(M, np.int32))
    nm = check(binding.load().orbx_search_local_points(
        device, kps.ctypes.data, desc.ctypes.data, n, _p(ur), bounds[0], bounds[1], bounds[2], bounds[3], sf.ctypes.data, len(sf),
        R.ctypes.data, t.ctypes.data, Ow.ctypes.data, cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["mbf"], log_sf,
        M, pos.ctypes.data, nrm.ctypes.data, dmin.ctypes.data, dmax.ctypes.data, md.ctypes.data, _p(cand),
        cos_limit, th, nnratio, int(far), th_far, assign.ctypes.data,
        *[out[k].ctypes.data for k in ("inview", "projx", "projy", "projxr", "depth", "viewcos", "level")]))
    return nm, assign, out


def SearchByProjectionSim3(kps, desc, bounds, scale_factors, mp: dict, th: float, ratio_hamming: float, matched=None, device: int = 0):
    """ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming) (src/ORBmatcher.cc:549, :681).
    mp: SoA dict desc,u,v,level,valid.  Returns (nmatches, assign)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps); sf = np.ascontiguousarray(scale_factors, np.float32)
    assign = np.full(n, -1, np.int32)
    held = None if matched is None else np.ascontiguousarray(matched, np.uint8)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, mu, mv, ml, mval = f("desc", np.uint8), f("u", np.float32), f("v", np.float32), f("level", np.int32), f("valid", np.uint8)
    nm = check(binding.load().orbx_search_by_projection_sim3(
        device, kps.ctypes.data, desc.ctypes.data, n, bounds[0], bounds[1], bounds[2], bounds[3], sf.ctypes.data, len(sf),
        len(mu), md.ctypes.data, mu.ctypes.data, mv.ctypes.data, ml.ctypes.data, mval.ctypes.data, th, ratio_hamming,
        _p(held), assign.ctypes.data))
    return nm, assign


def Fuse(kps, desc, uRight, bounds, scale_factors, inv_sigma2, mp: dict, th: float, device: int = 0):
    """The search of ORBmatcher::Fuse (src/ORBmatcher.cc:1823 with mp['ur'], :2051 without).  Returns (nFused, best_idx, best_dist)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps); sf = np.ascontiguousarray(scale_factors, np.float32); isg = np.ascontiguousarray(inv_sigma2, np.float32)
    ur = None if uRight is None else np.ascontiguousarray(uRight, np.float32)
    f = lambda k, t: np.ascontiguousarray(mp[k], t)
    md, mu, mv, ml, mval = f("desc", np.uint8), f("u", np.float32), f("v", np.float32), f("level", np.int32), f("valid", np.uint8)
    mur = f("ur", np.float32) if mp.get("ur") is not None else None
    bi = np.empty(len(mu), np.int32); bd = np.empty(len(mu), np.int32)
    nf = check(binding.load().orbx_fuse(
        device, kps.ctypes.data, desc.ctypes.data, n, _p(ur), bounds[0], bounds[1], bounds[2], bounds[3],
        sf.ctypes.data, isg.ctypes.data, len(sf), len(mu), md.ctypes.data, mu.ctypes.data, mv.ctypes.data, _p(mur),
        ml.ctypes.data, mval.ctypes.data, th, bi.ctypes.data, bd.ctypes.data))
    return nf, bi, bd


def SearchBySim3(kps1, desc1, kps2, desc2, bounds, scale_factors, rows1: dict, rows2: dict, th: float, device: int = 0):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:2201).  rows1: KF1's map points projected into KF2.  Returns (nFound, match12)."""
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); desc1 = np.ascontiguousarray(desc1, np.uint8)
    kps2 = np.ascontiguousarray(kps2, KP_DTYPE); desc2 = np.ascontiguousarray(desc2, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    g = lambda r, k, t: np.ascontiguousarray(r[k], t)
    a = [g(rows1, "desc", np.uint8), g(rows1, "u", np.float32), g(rows1, "v", np.float32), g(rows1, "level", np.int32), g(rows1, "valid", np.uint8)]
    b = [g(rows2, "desc", np.uint8), g(rows2, "u", np.float32), g(rows2, "v", np.float32), g(rows2, "level", np.int32), g(rows2, "valid", np.uint8)]
    m12 = np.full(len(kps1), -1, np.int32)
    nf = check(binding.load().orbx_search_by_sim3(
        device, kps1.ctypes.data, desc1.ctypes.data, len(kps1), kps2.ctypes.data, desc2.ctypes.data, len(kps2),
        bounds[0], bounds[1], bounds[2], bounds[3], sf.ctypes.data, len(sf),
        *[x.ctypes.data for x in a], *[x.ctypes.data for x in b], th, m12.ctypes.data))
    return nf, m12


def ComputeThreeMaxima(sizes):
    """ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2863) on the bin sizes."""
    sizes = np.ascontiguousarray(sizes, np.int32)
    out = np.zeros(3, np.int32)
    binding.load().orbx_three_maxima(sizes.ctypes.data, len(sizes), out[0:].ctypes.data, out[1:].ctypes.data, out[2:].ctypes.data)
    return tuple(int(v) for v in out)


# ---- two-camera rig (Frame::Nleft != -1) branches, callback-driven triangulation searches, fisheye stereo ----
VERDICT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int)


def _kf_args(k, flag_key):
    kps = np.ascontiguousarray(k["kps"], KP_DTYPE); desc = np.ascontiguousarray(k["desc"], np.uint8)
    flag = None if k.get(flag_key) is None else np.ascontiguousarray(k[flag_key], np.uint8)
    fn, fs, fi = (np.ascontiguousarray(k[x], np.int32) for x in ("fv_node", "fv_start", "fv_idx"))
    return kps, desc, flag, fn, fs, fi


def SearchByProjectionFrameRig(kpsL, kpsR, desc, bounds, scale_factors, last: dict, th: float, forward: bool, backward: bool,
                               check_orientation: bool = True, cur_held=None, device: int = 0):
    """ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono), CurrentFrame.Nleft != -1
    (src/ORBmatcher.cc:2469-2711).  last: SoA dict desc,u,v,ur,vr,invz,octave,angle,valid,obs.  Returns (nmatches, assign)."""
    kpsL = np.ascontiguousarray(kpsL, KP_DTYPE); kpsR = np.ascontiguousarray(kpsR, KP_DTYPE)
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kpsL) + len(kpsR)
    assign = np.full(n, -1, np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    held = None if cur_held is None else np.ascontiguousarray(cur_held, np.uint8)
    f = lambda k, t: np.ascontiguousarray(last[k], t)
    ld = f("desc", np.uint8)
    lu, lv, lur, lvr, lz = (f(k, np.float32) for k in ("u", "v", "ur", "vr", "invz"))
    lo, la, lval, lobs = f("octave", np.int32), f("angle", np.float32), f("valid", np.uint8), f("obs", np.uint8)
    nm = check(binding.load().orbx_search_by_projection_frame_rig(
        device, kpsL.ctypes.data, len(kpsL), kpsR.ctypes.data, len(kpsR), desc.ctypes.data, bounds[0], bounds[1], bounds[2], bounds[3],
        sf.ctypes.data, len(sf), len(lu), ld.ctypes.data, lu.ctypes.data, lv.ctypes.data, lur.ctypes.data, lvr.ctypes.data, lz.ctypes.data,
        lo.ctypes.data, la.ctypes.data, lval.ctypes.data, lobs.ctypes.data, th, int(forward), int(backward), int(check_orientation),
        _p(held), assign.ctypes.data))
    return nm, assign


def SearchByBoWRig(kf1: dict, kf2: dict, nleft2: int, nnratio: float, check_orientation: bool = True, device: int = 0):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) with F.Nleft != -1 (src/ORBmatcher.cc:314-547).  Returns (nmatches, matches[n2])."""
    a, b = _kf_args(kf1, "valid"), _kf_args(kf2, "valid")
    out = np.full(len(b[0]), -1, np.int32)
    nm = check(binding.load().orbx_search_by_bow_rig(
        device, a[0].ctypes.data, a[1].ctypes.data, len(a[0]), a[2].ctypes.data, a[3].ctypes.data, a[4].ctypes.data, a[5].ctypes.data, len(a[3]),
        b[0].ctypes.data, b[1].ctypes.data, len(b[0]), int(nleft2), b[3].ctypes.data, b[4].ctypes.data, b[5].ctypes.data, len(b[3]),
        nnratio, int(check_orientation), out.ctypes.data))
    return nm, out


def SearchForTriangulationCb(overload: int, kf1: dict, kf2: dict, verdict, only_stereo: bool = False, coarse: bool = False,
                             check_orientation: bool = True, device: int = 0):
    """SearchForTriangulation / SearchForTriangulation_ with a two-camera rig (overload 0 / 1) and the matchAndtriangulate
    overload (2), src/ORBmatcher.cc:1107 / :1388 / :1632.  verdict(idx1, idx2) -> bool is the camera model's test."""
    a, b = _kf_args(kf1, "has_mp"), _kf_args(kf2, "has_mp")
    m12 = np.full(len(a[0]), -1, np.int32)
    cb = VERDICT_FN(lambda user, i1, i2: int(bool(verdict(i1, i2))))
    args = []
    for k in (a, b):
        args += [k[0].ctypes.data, k[1].ctypes.data, len(k[0]), k[2].ctypes.data, k[3].ctypes.data, k[4].ctypes.data, k[5].ctypes.data, len(k[3])]
    nm = check(binding.load().orbx_search_for_triangulation_cb(device, int(overload), *args, int(only_stereo), int(coarse),
                                                               int(check_orientation), C.cast(cb, C.c_void_p), None, m12.ctypes.data))
    return nm, m12


def ComputeStereoFishEyeMatches(XL, XR, nL: int, monoLeft: int, nR: int, monoRight: int, frame: int = 0):
    """Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1440-1480) up to the triangulation gate, on the descriptors the two
    extractors left on the device.  Returns (descMatches, l2r[nL], r2l[nR], best_dist[nL])."""
    l2r = np.full(nL, -1, np.int32); r2l = np.full(nR, -1, np.int32); bd = np.full(nL, -1, np.int32)
    good = check(binding.load().orbx_stereo_fisheye_match(XL.h, XR.h, int(frame), int(nL), int(monoLeft), int(nR), int(monoRight),
                                                          l2r.ctypes.data, r2l.ctypes.data, bd.ctypes.data))
    return good, l2r, r2l, bd

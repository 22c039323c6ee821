"""Second source for the Frame.cc rows: oracle/_ref/libref_frame.so is the UNMODIFIED /root/reference/src/Frame.cc compiled with
the real include/Frame.h, the real ORBextractor.cc and the vendored DBoW2 (doubles only for the IMU / camera / MapPoint headers,
oracle/refshim_frame/).  The C restatements the GPU is compared with must equal what the reference's own member functions produce:
  Frame::ComputeStereoMatches           == orc_stereo_match          (mvuRight, mvDepth bit for bit)
  AssignFeaturesToGrid/GetFeaturesInArea == orc_grid_build / orc_grid_query (same indices in the same order)
  Frame::isInFrustum                    == orc_is_in_frustum         (every mTrack* field)
  Frame::ComputeStereoFishEyeMatches    == orc_stereo_fisheye_match
  Frame::Frame(imLeft, imRight, ...)    the whole rectified-stereo constructor (two ExtractORB threads -> ComputeStereoMatches)
"""
import ctypes as C
import pathlib

import numpy as np
import pytest

from test_gpu_search import SF, _frame, _scene

ROOT = pathlib.Path(__file__).resolve().parents[1]
VP = C.c_void_p


@pytest.fixture(scope="module")
def reff():
    path = ROOT / "oracle" / "_ref" / "libref_frame.so"
    if not path.exists():
        import ref_py
        try:
            ref_py.build()
        except Exception:
            pass
    if not path.exists():
        pytest.skip("oracle/_ref/libref_frame.so not built and /root/reference absent")
    L = C.CDLL(str(path))
    L.ref_frame_grid_build.restype = VP
    L.ref_frame_grid_build.argtypes = [VP, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]
    L.ref_frame_grid_destroy.argtypes = [VP]
    L.ref_frame_grid_query.argtypes = [VP, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, VP, C.c_int]
    L.ref_frame_stereo_matches.argtypes = [VP] * 6 + [C.c_int, VP, VP, C.c_int, VP, VP, C.c_int, VP, C.c_float, C.c_float, VP, VP]
    L.ref_frame_is_in_frustum.argtypes = [VP, VP, VP] + [C.c_float] * 10 + [C.c_int, C.c_int] + [VP] * 5 + [C.c_float] + [VP] * 7
    L.ref_frame_stereo_fisheye.argtypes = [VP, C.c_int, C.c_int, VP, C.c_int, C.c_int, VP, VP, VP]
    L.ref_frame_stereo_ctor.argtypes = ([VP, VP, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int] + [C.c_float] * 6
                                        + [VP, VP, C.POINTER(C.c_int), VP, VP, C.POINTER(C.c_int), C.c_int, VP, VP])
    return L


def _ref_stereo_matches(reff, EL, ER, kl, dl, kr, dr, mb, mbf):
    nlv = 8
    pl = [np.ascontiguousarray(EL.level_plane(l)) for l in range(nlv)]
    pr = [np.ascontiguousarray(ER.level_plane(l)) for l in range(nlv)]
    rows = np.array([p.shape[0] - 38 for p in pl], np.int32); cols = np.array([p.shape[1] - 38 for p in pl], np.int32)
    ptr = lambda ps: (C.c_void_p * nlv)(*[p.ctypes.data for p in ps])
    stp = lambda ps: (C.c_size_t * nlv)(*[p.strides[0] for p in ps])
    n = len(kl)
    uR = np.empty(n, np.float32); depth = np.empty(n, np.float32)
    sf = np.ascontiguousarray(EL.scale_factors, np.float32)
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr); dl = np.ascontiguousarray(dl); dr = np.ascontiguousarray(dr)
    reff.ref_frame_stereo_matches(ptr(pl), ptr(pr), rows.ctypes.data, cols.ctypes.data, stp(pl), stp(pr), nlv, sf.ctypes.data,
                                  kl.ctypes.data, n, dl.ctypes.data, kr.ctypes.data, len(kr), dr.ctypes.data, mb, mbf,
                                  uR.ctypes.data, depth.ctypes.data)
    return uR, depth


@pytest.mark.parametrize("cfg,seed,max_disp", [("C2", 11, 40), ("C2", 12, 90), ("C4", 21, 40), ("C3", 31, 25)])
def test_compute_stereo_matches_equals_reference(orc, oracle, pkg_synth, reff, cfg, seed, max_disp):
    c = pkg_synth.CONFIGS[cfg]
    rows, cols, nf = c["rows"], c["cols"], c["nfeatures"]
    mbf = c["mbf"]; mb = mbf / c["fx"]
    L, R = pkg_synth.synth_stereo_pair(rows, cols, seed, max_disp=max_disp)[:2]
    EL, ER = orc.Extractor(nf, oracle=oracle), orc.Extractor(nf, oracle=oracle)
    _, kl, dl = EL(L, (0, 0)); _, kr, dr = ER(R, (0, 0))
    u_o, d_o, _, _ = orc.stereo_match(EL, ER, kl, dl, kr, dr, mb, mbf)
    u_r, d_r = _ref_stereo_matches(reff, EL, ER, kl, dl, kr, dr, mb, mbf)
    assert (u_o >= 0).sum() > 80
    assert np.array_equal(u_o.view(np.uint32), u_r.view(np.uint32)), f"{np.sum(u_o != u_r)} mvuRight differ"
    assert np.array_equal(d_o.view(np.uint32), d_r.view(np.uint32))
    assert ((u_o < 0) & (d_o < 0)).any() and (u_o >= 0).any()


def test_compute_stereo_matches_identical_eyes(orc, oracle, pkg_synth, reff):
    """zero disparity everywhere: the `disparity <= 0 -> 0.01` clamp (src/Frame.cc:1246-1250) and the sub-pixel parabola's 0/0"""
    c = pkg_synth.CONFIGS["C2"]
    img = pkg_synth.synth_frame(c["rows"], c["cols"], 5)
    E1, E2 = orc.Extractor(800, oracle=oracle), orc.Extractor(800, oracle=oracle)
    _, k1, d1 = E1(img, (0, 0)); _, k2, d2 = E2(img, (0, 0))
    mbf = c["mbf"]; mb = mbf / c["fx"]
    u_o, d_o, _, _ = orc.stereo_match(E1, E2, k1, d1, k2, d2, mb, mbf)
    u_r, d_r = _ref_stereo_matches(reff, E1, E2, k1, d1, k2, d2, mb, mbf)
    assert np.array_equal(u_o.view(np.uint32), u_r.view(np.uint32)) and np.array_equal(d_o.view(np.uint32), d_r.view(np.uint32))


@pytest.mark.parametrize("n,rows,cols,bounds,seed", [(1500, 480, 752, None, 1), (4000, 376, 1241, None, 2),
                                                     (1200, 480, 752, (-21.7, 771.3, -14.2, 497.9), 3), (50, 480, 752, None, 4)])
def test_frame_grid_equals_reference(orc, oracle, reff, n, rows, cols, bounds, seed):
    """AssignFeaturesToGrid + PosInGrid + GetFeaturesInArea (src/Frame.cc:446-480, 784-899) against orc_grid_*: same keypoint
    indices in the same order, with and without the level filters, windows hanging over every border."""
    rng = np.random.default_rng(seed)
    kps, _, _, _ = _frame(orc, rng, n, rows, cols)
    b = bounds or (0.0, float(cols), 0.0, float(rows))
    if bounds:                                   # undistorted keypoints may fall outside the image; some outside the bounds too
        kps["x"] += rng.normal(0, 8, n).astype(np.float32); kps["y"] += rng.normal(0, 8, n).astype(np.float32)
    g = orc.Grid(oracle, kps, *b)
    F = reff.ref_frame_grid_build(kps.ctypes.data, n, *[C.c_float(v) for v in b])
    out = np.empty(n + 1, np.int32)
    hits = 0
    for q in range(1500):
        x = float(np.float32(rng.uniform(b[0] - 40, b[1] + 40))); y = float(np.float32(rng.uniform(b[2] - 40, b[3] + 40)))
        r = float(np.float32(rng.choice([1.0, 4.0, 12.0, 35.0, 120.0])))
        if q % 3 == 0:
            lo, hi = -1, -1
        elif q % 3 == 1:
            lo = int(rng.integers(0, 8)); hi = int(rng.integers(lo, 8))
        else:
            lo, hi = int(rng.integers(0, 8)), -1
        want = g.query(x, y, r, lo, hi)
        cnt = reff.ref_frame_grid_query(F, x, y, r, lo, hi, out.ctypes.data, n + 1)
        assert cnt == len(want) and np.array_equal(out[:cnt], want), (q, x, y, r, lo, hi)
        hits += cnt
    assert hits > 1000 or n < 100
    reff.ref_frame_grid_destroy(F)


@pytest.mark.parametrize("seed,cos_limit", [(91, 0.5), (92, 0.5), (93, 0.8)])
def test_is_in_frustum_equals_reference(orc, oracle, reff, seed, cos_limit):
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1200
    kps, desc, _, _ = _frame(orc, rng, n, rows, cols)
    cam = dict(fx=435.2047, fy=435.2047, cx=367.215, cy=252.375, mbf=47.90639)
    pose, pts = _scene(rng, kps, desc, rows, cols, cam, 4000)
    bounds = (0.0, float(cols), 0.0, float(rows))
    log_sf = float(np.log(np.float32(1.2)))
    n_o, o = orc.is_in_frustum(oracle, pose, cam, bounds, log_sf, 8, pts, cos_limit)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    R, t, Ow = f32(pose["Rcw"]).reshape(9), f32(pose["tcw"]), f32(pose["Ow"])
    pos, nrm, dmin, dmax = f32(pts["pos"]), f32(pts["normal"]), f32(pts["min_dist"]), f32(pts["max_dist"])
    cand = np.ascontiguousarray(pts["candidate"], np.uint8)
    m = len(dmin)
    r = dict(inview=np.empty(m, np.uint8), projx=np.empty(m, np.float32), projy=np.empty(m, np.float32), projxr=np.empty(m, np.float32),
             depth=np.empty(m, np.float32), viewcos=np.empty(m, np.float32), level=np.empty(m, np.int32))
    n_r = reff.ref_frame_is_in_frustum(R.ctypes.data, t.ctypes.data, Ow.ctypes.data, cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["mbf"],
                                       *bounds, log_sf, 8, m, pos.ctypes.data, nrm.ctypes.data, dmin.ctypes.data, dmax.ctypes.data,
                                       cand.ctypes.data, cos_limit, *[r[k].ctypes.data for k in ("inview", "projx", "projy", "projxr", "depth", "viewcos", "level")])
    if n_r <= -1000:
        # the reference derives Ow = -Rcw^T tcw itself (UpdatePoseMatrices, cv::Mat product = double accumulation); the scene's Ow is
        # the same expression in numpy, so they normally agree - if not, the comparison below is on the reference's Ow
        n_r = -1000 - n_r
    d = np.flatnonzero(o["inview"] != r["inview"])
    assert len(d) == 0, (len(d), {k: (o[k][d[0]], r[k][d[0]]) for k in o}, pos[d[0]], dmin[d[0]], dmax[d[0]], cand[d[0]])
    assert n_o == n_r > 300
    v = o["inview"] != 0
    for k in ("projx", "projy", "projxr", "depth", "viewcos"):
        assert np.array_equal(o[k][v].view(np.uint32), r[k][v].view(np.uint32)), k
    assert np.array_equal(o["level"][v], r["level"][v])


@pytest.mark.parametrize("nL,nR,monoL,monoR,seed", [(900, 800, 300, 250, 1), (500, 700, 0, 0, 2), (300, 2, 10, 0, 3), (200, 1, 0, 0, 4), (50, 60, 50, 10, 5)])
def test_stereo_fisheye_matches_equal_reference(orc, oracle, reff, nL, nR, monoL, monoR, seed):
    rng = np.random.default_rng(seed)
    dR = rng.integers(0, 256, (nR, 32), dtype=np.uint8)
    dL = rng.integers(0, 256, (nL, 32), dtype=np.uint8)
    for i in range(monoL, nL):                                   # many left rows are near-copies of a right row (ratio test passes)
        if nR > monoR and rng.random() < 0.6:
            j = int(rng.integers(monoR, nR))
            dL[i] = dR[j]
            for b in rng.integers(0, 256, int(rng.integers(0, 12))):
                dL[i, b >> 3] ^= 1 << (b & 7)
    good_o, l2r_o, r2l_o, _ = orc.stereo_fisheye_match(oracle, dL, monoL, dR, monoR)
    l2r = np.empty(nL, np.int32); r2l = np.empty(nR, np.int32); bd = np.empty(nL, np.int32)
    good_r = reff.ref_frame_stereo_fisheye(dL.ctypes.data, nL, monoL, dR.ctypes.data, nR, monoR, l2r.ctypes.data, r2l.ctypes.data, bd.ctypes.data)
    assert good_o == good_r
    assert np.array_equal(l2r_o, l2r) and np.array_equal(r2l_o, r2l)
    if nR - monoR >= 2 and nL - monoL > 100:
        assert good_o > 20


@pytest.mark.parametrize("cfg,seed", [("C2", 41), ("C4", 42)])
def test_stereo_constructor_end_to_end(orc, oracle, pkg_synth, reff, cfg, seed):
    """Frame::Frame(imLeft, imRight, ...) (src/Frame.cc:103-225) run as the reference runs it: two ExtractORB threads, then
    ComputeStereoMatches on mvImagePyramid.  The two extractions use the reference's literal (heap-address) quadtree tie-break, so
    a handful of keypoints differ from the canonical oracle's; the pyramids do not depend on it - so the oracle's stereo matcher,
    fed with the constructor's own keypoints and descriptors, must reproduce the constructor's mvuRight / mvDepth exactly."""
    c = pkg_synth.CONFIGS[cfg]
    rows, cols, nf = c["rows"], c["cols"], c["nfeatures"]
    mbf = c["mbf"]
    L, R = pkg_synth.synth_stereo_pair(rows, cols, seed)[:2]
    L = np.ascontiguousarray(L); R = np.ascontiguousarray(R)
    cap = 4 * nf
    kl = np.empty(cap, orc.KP_DTYPE); kr = np.empty(cap, orc.KP_DTYPE)
    dl = np.empty((cap, 32), np.uint8); dr = np.empty((cap, 32), np.uint8)
    uR = np.empty(cap, np.float32); depth = np.empty(cap, np.float32)
    nL, nR = C.c_int(), C.c_int()
    rc = reff.ref_frame_stereo_ctor(L.ctypes.data, R.ctypes.data, rows, cols, L.strides[0], nf, 1.2, 8, 20, 7,
                                    c["fx"], c["fx"], cols / 2.0, rows / 2.0, mbf, 35.0,
                                    kl.ctypes.data, dl.ctypes.data, C.byref(nL), kr.ctypes.data, dr.ctypes.data, C.byref(nR), cap,
                                    uR.ctypes.data, depth.ctypes.data)
    assert rc == 0
    nL, nR = nL.value, nR.value
    EL, ER = orc.Extractor(nf, oracle=oracle), orc.Extractor(nf, oracle=oracle)
    _, kl_o, dl_o = EL(L, (0, 0)); _, kr_o, dr_o = ER(R, (0, 0))
    assert abs(nL - len(kl_o)) <= 8 and abs(nR - len(kr_o)) <= 8
    same = len(set(map(bytes, kl[:nL])) & set(map(bytes, kl_o)))
    assert same >= 0.97 * nL                                   # literal vs canonical tie-break: < 3 % of the keypoints move
    # reference quirk: this constructor runs ComputeStereoMatches (:137) before it assigns mb = mbf / fx (:163); the wrapper builds
    # the Frame in zero-filled storage, so minZ = mb = 0 and maxD = mbf / 0 = +inf there (no upper disparity bound)
    u_o, d_o, _, _ = orc.stereo_match(EL, ER, kl[:nL], dl[:nL], kr[:nR], dr[:nR], 0.0, mbf)
    assert (u_o >= 0).sum() > 80
    bad = np.flatnonzero(u_o != uR[:nL])
    assert len(bad) == 0, [(int(i), float(u_o[i]), float(uR[i]), float(d_o[i]), float(depth[i]), kl[i]) for i in bad[:5]]
    assert np.array_equal(d_o.view(np.uint32), depth[:nL].view(np.uint32))

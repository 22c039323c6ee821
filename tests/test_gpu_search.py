"""GPU parity (through the C ABI) for the guided searches of src/ORBmatcher.cc:
GetFeaturesInArea windows, SearchByProjection(Frame,Frame), SearchForInitialization and
SearchForTriangulation[_], each against the oracle restatement on the same seeded inputs.  Bar: identical
match vectors and counts (integer work; the float gates are evaluated operation by operation)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SF = np.cumprod(np.concatenate([[1.0], np.full(7, 1.2)])).astype(np.float32)
SIGMA2 = (SF * SF).astype(np.float32)


def _flip(rng, d, lo, hi):
    bits = np.unpackbits(d, axis=1)
    for i in range(len(d)):
        k = int(rng.integers(lo, hi + 1))
        bits[i, rng.choice(256, k, replace=False)] ^= 1
    return np.packbits(bits, axis=1)


def _frame(pkg, rng, n, rows, cols, pool=None, quant=1.0):
    """n random keypoints; descriptors drawn from a small pool (+ a few flipped bits) so that equal
    distances and repeated best candidates occur."""
    kps = np.zeros(n, pkg.KP_DTYPE)
    kps["x"] = np.round(rng.uniform(0, cols, n) / quant) * quant
    kps["y"] = np.round(rng.uniform(0, rows, n) / quant) * quant
    kps["octave"] = rng.choice(8, n, p=[.3, .22, .16, .12, .08, .06, .04, .02])
    kps["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    kps["size"] = 31.0 * SF[kps["octave"]]
    kps["class_id"] = -1
    if pool is None:
        pool = rng.integers(0, 256, (max(8, n // 6), 32), dtype=np.uint8)
    src = rng.integers(0, len(pool), n)
    desc = _flip(rng, pool[src].copy(), 0, 12)
    return kps, desc, pool, src


def _moved(pkg, rng, kps, desc, rows, cols, shift=(4.0, -3.0), jitter=2.0, flips=(0, 40)):
    """A second view of the same points: shifted and jittered positions, noisy descriptors/angles."""
    n = len(kps)
    k2 = kps.copy()
    k2["x"] = np.clip(kps["x"] + shift[0] + rng.normal(0, jitter, n), 0, cols - 1).astype(np.float32)
    k2["y"] = np.clip(kps["y"] + shift[1] + rng.normal(0, jitter, n), 0, rows - 1).astype(np.float32)
    k2["octave"] = np.clip(kps["octave"] + rng.choice([-1, 0, 0, 0, 1], n), 0, 7)
    ang = kps["angle"] + np.where(rng.random(n) < 0.8, rng.normal(0, 4, n), rng.uniform(0, 360, n))
    k2["angle"] = np.mod(ang, 360).astype(np.float32)
    d2 = _flip(rng, desc.copy(), *flips)
    perm = rng.permutation(n)
    return k2[perm], d2[perm], perm


@pytest.mark.parametrize("n,nq,seed", [(1200, 900, 1), (3000, 64, 2), (40, 500, 3), (0, 10, 4)])
def test_window_candidates(pkg, orc, oracle, n, nq, seed):
    rng = np.random.default_rng(seed)
    rows, cols = 480, 752
    kps, desc, _, _ = _frame(pkg, rng, n, rows, cols)
    qd = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    qx = rng.uniform(-40, cols + 40, nq).astype(np.float32); qy = rng.uniform(-40, rows + 40, nq).astype(np.float32)
    qr = rng.choice([3.0, 7.5, 15.0, 100.0, -1.0, 0.0, 2000.0], nq).astype(np.float32)
    qmin = rng.integers(-1, 6, nq).astype(np.int32)
    qmax = np.where(rng.random(nq) < 0.3, -1, qmin + rng.integers(0, 3, nq)).astype(np.int32)
    start, idx, dist = pkg.GetFeaturesInArea(kps, desc, (0.0, float(cols), 0.0, float(rows)), qd, qx, qy, qr, qmin, qmax)
    g = orc.Grid(oracle, kps, 0.0, float(cols), 0.0, float(rows))
    tot = 0
    for q in range(nq):
        ref = g.query(float(qx[q]), float(qy[q]), float(qr[q]), int(qmin[q]), int(qmax[q])) if qr[q] >= 0 else np.zeros(0, np.int32)
        got = idx[start[q]:start[q + 1]]
        assert np.array_equal(got, ref), f"query {q}"
        for k, i2 in enumerate(ref):
            assert dist[start[q] + k] == oracle.hamming(qd[q], desc[i2])
        tot += len(ref)
    assert start[nq] == tot
    if n >= 1000:
        assert tot > 1000


def test_window_candidates_capacity(pkg):
    rng = np.random.default_rng(0)
    kps, desc, _, _ = _frame(pkg, rng, 500, 480, 752)
    one = np.ones(4, np.float32)
    with pytest.raises(pkg.OrbxError):
        pkg.GetFeaturesInArea(kps, desc, (0.0, 752.0, 0.0, 480.0), desc[:4], 300 * one, 200 * one, 500 * one,
                              -np.ones(4, np.int32), -np.ones(4, np.int32), cap=16)


@pytest.mark.parametrize("stereo,direction,orient,seed", [(True, "none", True, 11), (True, "forward", True, 12),
                                                          (True, "backward", False, 13), (False, "none", True, 14)])
def test_search_by_projection_frame(pkg, orc, oracle, stereo, direction, orient, seed):
    """TrackWithMotionModel's matcher (th = 15 mono / 7 stereo, src/Tracking.cc)."""
    rng = np.random.default_rng(seed)
    rows, cols = 480, 752
    n = 1500
    lkps, ldesc, _, _ = _frame(pkg, rng, n, rows, cols)
    ckps, cdesc, perm = _moved(pkg, rng, lkps, ldesc, rows, cols)
    uR = None
    if stereo:
        uR = np.where(rng.random(n) < 0.7, ckps["x"] - rng.uniform(1, 40, n), -1).astype(np.float32)
    invz = np.where(rng.random(n) < 0.97, rng.uniform(0.02, 0.5, n), -0.1).astype(np.float32)
    last = dict(desc=_flip(rng, ldesc.copy(), 0, 10),
                u=(lkps["x"] + 4.0 + rng.normal(0, 3, n)).astype(np.float32),
                v=(lkps["y"] - 3.0 + rng.normal(0, 3, n)).astype(np.float32),
                invz=invz, octave=lkps["octave"].astype(np.int32), angle=lkps["angle"],
                valid=(rng.random(n) < 0.85).astype(np.uint8), obs=(rng.random(n) < 0.9).astype(np.uint8))
    held = (rng.random(n) < 0.03).astype(np.uint8)
    th, mbf = (7.0, 40.0) if stereo else (15.0, 0.0)
    fwd, bwd = direction == "forward", direction == "backward"
    bounds = (0.0, float(cols), 0.0, float(rows))
    g = orc.Grid(oracle, ckps, *bounds)
    nm_o, as_o = orc.search_by_projection_frame(g, cdesc, uR, SF, last, th, mbf, fwd, bwd, orient, held)
    nm_g, as_g = pkg.SearchByProjectionFrame(ckps, cdesc, uR, bounds, SF, last, th, mbf, fwd, bwd, orient, held)
    assert nm_o > 150
    assert nm_g == nm_o
    assert np.array_equal(as_g, as_o), f"{np.sum(as_g != as_o)} assignments differ"


@pytest.mark.parametrize("window,orient,seed", [(100, True, 21), (100, False, 22), (30, True, 23)])
def test_search_for_initialization(pkg, orc, oracle, window, orient, seed):
    rng = np.random.default_rng(seed)
    rows, cols = 480, 752
    n = 2000
    k1, d1, _, _ = _frame(pkg, rng, n, rows, cols, quant=1.0)
    k2, d2, perm = _moved(pkg, rng, k1, d1, rows, cols, shift=(12.0, 5.0), flips=(0, 30))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    bounds = (0.0, float(cols), 0.0, float(rows))
    g2 = orc.Grid(oracle, k2, *bounds)
    nm_o, m_o, pm_o = orc.search_for_initialization(oracle, k1, d1, g2, d2, prev, window, 0.9, orient)
    nm_g, m_g, pm_g = pkg.SearchForInitialization(k1, d1, k2, d2, bounds, prev, window, 0.9, orient)
    assert nm_o > 50
    assert nm_g == nm_o == int(np.sum(m_o >= 0))
    assert np.array_equal(m_g, m_o)
    assert np.array_equal(pm_g, pm_o)
    # second round feeds the updated vbPrevMatched back, as Tracking::MonocularInitialization does
    nm_o2, m_o2, pm_o2 = orc.search_for_initialization(oracle, k1, d1, g2, d2, pm_o, window, 0.9, orient)
    nm_g2, m_g2, pm_g2 = pkg.SearchForInitialization(k1, d1, k2, d2, bounds, pm_g, window, 0.9, orient)
    assert nm_g2 == nm_o2 and np.array_equal(m_g2, m_o2) and np.array_equal(pm_g2, pm_o2)


def _feature_vector(rng, node_of, drop=0.0):
    """CSR FeatureVector: node id -> keypoint indices in ascending index order (DBoW2 push_back order)."""
    keep = rng.random(len(node_of)) >= drop
    nodes = np.unique(node_of[keep])
    start = [0]; idx = []
    for nd in nodes:
        m = np.nonzero((node_of == nd) & keep)[0]
        idx.extend(m.tolist()); start.append(len(idx))
    return nodes.astype(np.int32), np.array(start, np.int32), np.array(idx, np.int32)


@pytest.mark.parametrize("variant,only_stereo,coarse,stereo,seed", [(0, False, False, False, 31), (0, False, False, True, 32),
                                                                  (1, False, False, True, 33), (0, True, False, True, 34),
                                                                  (0, False, True, False, 35), (1, False, True, False, 36)])
def test_search_for_triangulation(pkg, orc, oracle, variant, only_stereo, coarse, stereo, seed):
    rng = np.random.default_rng(seed)
    rows, cols = 480, 752
    n = 1800
    k1, d1, pool, src = _frame(pkg, rng, n, rows, cols)
    # second keyframe: pure sideways translation -> horizontal epipolar lines, F12 = [t]x up to K
    k2, d2, perm = _moved(pkg, rng, k1, d1, rows, cols, shift=(-18.0, 0.0), jitter=0.8, flips=(0, 45))
    fx = fy = 458.0; cx, cy = 367.0, 248.0
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    t12 = np.array([0.11, 0.002, -0.004]); tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    F12 = (np.linalg.inv(K).T @ tx @ np.eye(3) @ np.linalg.inv(K)).astype(np.float32)
    ep = np.array([cx + 120.0, cy - 15.0], np.float32)            # an epipole inside the image exercises :1241-1250
    node1 = (src % 97).astype(np.int32); node2 = node1[perm]
    wrong = rng.random(n) < 0.05
    node2 = np.where(wrong, rng.integers(0, 97, n), node2).astype(np.int32)
    kf1 = dict(kps=k1, desc=d1, has_mp=(rng.random(n) < 0.35).astype(np.uint8), uRight=None)
    kf2 = dict(kps=k2, desc=d2, has_mp=(rng.random(n) < 0.35).astype(np.uint8), uRight=None)
    if stereo:
        kf1["uRight"] = np.where(rng.random(n) < 0.5, k1["x"] - 10, -1).astype(np.float32)
        kf2["uRight"] = np.where(rng.random(n) < 0.5, k2["x"] - 10, -1).astype(np.float32)
    kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = _feature_vector(rng, node1, 0.02)
    kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = _feature_vector(rng, node2, 0.02)
    nm_o, m_o = orc.search_for_triangulation(oracle, kf1, kf2, SF, SIGMA2, ep, F12, only_stereo, coarse, True, variant)
    nm_g, m_g = pkg.SearchForTriangulation(kf1, kf2, SF, SIGMA2, ep, F12, only_stereo, coarse, True, variant)
    assert nm_o > (5 if only_stereo else 40), nm_o
    assert nm_g == nm_o
    assert np.array_equal(m_g, m_o), f"{np.sum(m_g != m_o)} matches differ"
    nm_o2, m_o2 = orc.search_for_triangulation(oracle, kf1, kf2, SF, SIGMA2, ep, F12, only_stereo, coarse, False, variant)
    nm_g2, m_g2 = pkg.SearchForTriangulation(kf1, kf2, SF, SIGMA2, ep, F12, only_stereo, coarse, False, variant)
    assert nm_g2 == nm_o2 and np.array_equal(m_g2, m_o2)


def test_search_empty_inputs(pkg):
    z = np.zeros(0, pkg.KP_DTYPE); zd = np.zeros((0, 32), np.uint8)
    rng = np.random.default_rng(1)
    k, d, _, _ = _frame(pkg, rng, 50, 480, 752)
    b = (0.0, 752.0, 0.0, 480.0)
    last = dict(desc=zd, u=[], v=[], invz=[], octave=[], angle=[], valid=[], obs=[])
    assert pkg.SearchByProjectionFrame(k, d, None, b, SF, last, 15.0, 0.0, False, False)[0] == 0
    assert pkg.SearchForInitialization(z, zd, k, d, b, np.zeros((0, 2), np.float32))[0] == 0
    nm, m12, _ = pkg.SearchForInitialization(k, d, z, zd, b, np.zeros((50, 2), np.float32))
    assert nm == 0 and np.all(m12 == -1)


@pytest.mark.parametrize("mode,nnratio,orient,seed", [(0, 0.7, True, 41), (0, 0.9, False, 42), (1, 0.8, True, 43), (1, 0.6, True, 44)])
def test_search_by_bow(pkg, orc, oracle, mode, nnratio, orient, seed):
    """TrackReferenceKeyFrame / Relocalization (mode 0: matcher(0.7|0.75,true)) and the loop-closing KF-KF search (mode 1)."""
    rng = np.random.default_rng(seed)
    rows, cols = 480, 752
    n = 1600
    k1, d1, pool, src = _frame(pkg, rng, n, rows, cols)
    k2, d2, perm = _moved(pkg, rng, k1, d1, rows, cols, flips=(0, 50))
    node1 = (src % 89).astype(np.int32); node2 = node1[perm]
    node2 = np.where(rng.random(n) < 0.05, rng.integers(0, 89, n), node2).astype(np.int32)
    kf1 = dict(kps=k1, desc=d1, valid=(rng.random(n) < 0.7).astype(np.uint8))
    kf2 = dict(kps=k2, desc=d2, valid=None if mode == 0 else (rng.random(n) < 0.7).astype(np.uint8))
    kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = _feature_vector(rng, node1, 0.02)
    kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = _feature_vector(rng, node2, 0.02)
    nm_o, m_o = orc.search_by_bow(oracle, mode, kf1, kf2, nnratio, orient)
    nm_g, m_g = pkg.SearchByBoW(mode, kf1, kf2, nnratio, orient)
    assert nm_o > 40, nm_o
    assert nm_g == nm_o == int(np.sum(m_o >= 0))
    assert np.array_equal(m_g, m_o), f"{np.sum(m_g != m_o)} matches differ"


@pytest.mark.parametrize("th,orb_dist,orient,seed", [(10.0, 100, True, 51), (3.0, 64, True, 52), (10.0, 100, False, 53)])
def test_search_by_projection_keyframe(pkg, orc, oracle, th, orb_dist, orient, seed):
    """Tracking::Relocalization: matcher2(0.9,true).SearchByProjection(mCurrentFrame, vpCandidateKFs[i], sFound, 10, 100) / (…, 3, 64)."""
    rng = np.random.default_rng(seed)
    rows, cols = 480, 752
    n = 1500
    kkps, kdesc, _, _ = _frame(pkg, rng, n, rows, cols)
    ckps, cdesc, perm = _moved(pkg, rng, kkps, kdesc, rows, cols)
    mp = dict(desc=_flip(rng, kdesc.copy(), 0, 10),
              u=(kkps["x"] + 4.0 + rng.normal(0, 2, n)).astype(np.float32), v=(kkps["y"] - 3.0 + rng.normal(0, 2, n)).astype(np.float32),
              level=np.clip(kkps["octave"] + rng.choice([-1, 0, 0, 1], n), 0, 7).astype(np.int32), angle=kkps["angle"],
              valid=(rng.random(n) < 0.8).astype(np.uint8))
    held = (rng.random(n) < 0.1).astype(np.uint8)
    bounds = (0.0, float(cols), 0.0, float(rows))
    g = orc.Grid(oracle, ckps, *bounds)
    nm_o, as_o = orc.search_by_projection_keyframe(g, cdesc, SF, mp, th, orb_dist, orient, held)
    nm_g, as_g = pkg.SearchByProjectionKeyFrame(ckps, cdesc, bounds, SF, mp, th, orb_dist, orient, held)
    assert nm_o > 100, nm_o
    assert nm_g == nm_o
    assert np.array_equal(as_g, as_o), f"{np.sum(as_g != as_o)} assignments differ"


def _projected_rows(pkg, rng, kkps, kdesc, n, rows, cols, with_ur=False):
    """Map-point rows projected near the keypoints they came from (some far away / outside the image / invalid)."""
    mp = dict(desc=_flip(rng, kdesc.copy(), 0, 30),
              u=(kkps["x"] + 3.0 + rng.normal(0, 2.5, n)).astype(np.float32), v=(kkps["y"] - 2.0 + rng.normal(0, 2.5, n)).astype(np.float32),
              level=np.clip(kkps["octave"] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32),
              valid=(rng.random(n) < 0.85).astype(np.uint8))
    far = rng.random(n) < 0.1
    mp["u"][far] = rng.uniform(-30, cols + 30, int(far.sum())).astype(np.float32)
    mp["u"][:3] = [0.0, float(cols), -0.5]              # IsInImage: x >= min && x < max
    mp["v"][3:6] = [0.0, float(rows), float(rows) - 0.25]
    if with_ur:
        mp["ur"] = (mp["u"] - rng.uniform(0, 40, n)).astype(np.float32)
    return mp


@pytest.mark.parametrize("th,ratio,seed,use_held", [(3.0, 1.0, 61, True), (8.0, 1.5, 62, True), (5.0, 0.6, 63, False), (8.0, 6.0, 64, True)])
def test_search_by_projection_sim3(pkg, orc, oracle, th, ratio, seed, use_held):
    """LoopClosing: matcher.SearchByProjection(mpCurrentKF, mScw, vpPoints, vpMatched, th, ratioHamming) (:549 and :681)."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1500
    kkps, kdesc, _, _ = _frame(pkg, rng, n, rows, cols)
    ckps, cdesc, perm = _moved(pkg, rng, kkps, kdesc, rows, cols)
    mp = _projected_rows(pkg, rng, kkps, kdesc, n, rows, cols)
    held = (rng.random(n) < 0.15).astype(np.uint8) if use_held else None
    bounds = (0.0, float(cols), 0.0, float(rows))
    g = orc.Grid(oracle, ckps, *bounds)
    nm_o, as_o = orc.search_by_projection_sim3(g, cdesc, SF, mp, th, ratio, held)
    nm_g, as_g = pkg.SearchByProjectionSim3(ckps, cdesc, bounds, SF, mp, th, ratio, held)
    assert nm_o > 50, nm_o
    assert nm_g == nm_o
    assert np.array_equal(as_g, as_o), f"{np.sum(as_g != as_o)} assignments differ"
    if use_held:
        assert not np.any((as_g >= 0) & (held != 0))


@pytest.mark.parametrize("form,stereo,th,seed", [("pose", True, 3.0, 71), ("pose", False, 3.0, 72), ("sim3", False, 4.0, 73), ("pose", True, 10.0, 74)])
def test_fuse(pkg, orc, oracle, form, stereo, th, seed):
    """LocalMapping::SearchInNeighbors: matcher.Fuse(pKFi, vpMapPointMatches) (:1823); LoopClosing::SearchAndFuse: Fuse(pKF, Scw, ...) (:2051)."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1800
    kkps, kdesc, _, _ = _frame(pkg, rng, n, rows, cols, quant=0.5)
    ckps, cdesc, perm = _moved(pkg, rng, kkps, kdesc, rows, cols)
    mp = _projected_rows(pkg, rng, kkps, kdesc, n, rows, cols, with_ur=(form == "pose"))
    uR = None
    if stereo:
        uR = np.where(rng.random(n) < 0.6, ckps["x"] - rng.uniform(0, 40, n), -1.0).astype(np.float32)
    bounds = (0.0, float(cols), 0.0, float(rows))
    isig2 = (1.0 / SIGMA2).astype(np.float32)
    g = orc.Grid(oracle, ckps, *bounds)
    nf_o, bi_o, bd_o = orc.fuse(g, cdesc, uR, SF, isig2, mp, th)
    nf_g, bi_g, bd_g = pkg.Fuse(ckps, cdesc, uR, bounds, SF, isig2, mp, th)
    assert nf_o > 30, nf_o
    assert nf_g == nf_o
    assert np.array_equal(bi_g, bi_o), f"{np.sum(bi_g != bi_o)} fuse targets differ"
    assert np.array_equal(bd_g, bd_o)


@pytest.mark.parametrize("th,seed", [(7.5, 81), (3.0, 82)])
def test_search_by_sim3(pkg, orc, oracle, th, seed):
    """LoopClosing::ComputeSim3: matcher.SearchBySim3(mpCurrentKF, pKF, vpMapPointMatches, s, R, t, 7.5) (:2201)."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1400
    k1, d1, _, _ = _frame(pkg, rng, n, rows, cols)
    k2, d2, perm = _moved(pkg, rng, k1, d1, rows, cols, flips=(0, 25))
    inv = np.empty(n, np.int64); inv[perm] = np.arange(n)          # k2[inv[i]] is the moved copy of k1[i]
    # rows 1: KF1's map points (descriptor ~ d1) projected where the point sits in KF2, and vice versa
    rows1 = dict(desc=_flip(rng, d1.copy(), 0, 10), u=(k2["x"][inv] + rng.normal(0, 1.5, n)).astype(np.float32),
                 v=(k2["y"][inv] + rng.normal(0, 1.5, n)).astype(np.float32),
                 level=np.clip(k2["octave"][inv] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32), valid=(rng.random(n) < 0.8).astype(np.uint8))
    rows2 = dict(desc=_flip(rng, d2.copy(), 0, 10), u=(k1["x"][perm] + rng.normal(0, 1.5, n)).astype(np.float32),
                 v=(k1["y"][perm] + rng.normal(0, 1.5, n)).astype(np.float32),
                 level=np.clip(k1["octave"][perm] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32), valid=(rng.random(n) < 0.8).astype(np.uint8))
    bounds = (0.0, float(cols), 0.0, float(rows))
    g1, g2 = orc.Grid(oracle, k1, *bounds), orc.Grid(oracle, k2, *bounds)
    nf_o, m_o = orc.search_by_sim3(g1, d1, g2, d2, SF, rows1, rows2, th)
    nf_g, m_g = pkg.SearchBySim3(k1, d1, k2, d2, bounds, SF, rows1, rows2, th)
    assert nf_o > 100, nf_o
    assert nf_g == nf_o
    assert np.array_equal(m_g, m_o), f"{np.sum(m_g != m_o)} matches differ"


def test_window_best_empty_inputs(pkg):
    kps = np.zeros(0, pkg.KP_DTYPE); desc = np.zeros((0, 32), np.uint8)
    mp = dict(desc=np.zeros((4, 32), np.uint8), u=np.ones(4, np.float32), v=np.ones(4, np.float32), level=np.zeros(4, np.int32), valid=np.ones(4, np.uint8))
    bounds = (0.0, 100.0, 0.0, 100.0)
    assert pkg.Fuse(kps, desc, None, bounds, SF, (1 / SIGMA2), mp, 3.0)[0] == 0
    assert pkg.SearchByProjectionSim3(kps, desc, bounds, SF, mp, 3.0, 1.0)[0] == 0


def _scene(rng, kps, desc, rows, cols, cam, nmp, nlevels=8):
    """A camera pose and nmp world points: 70 % back-projected from keypoints of the frame at random depths (so that
    they project onto features), the rest scattered around / behind the camera.  Normals, distance-invariance
    ranges and descriptors as MapPoint::UpdateNormalAndDepth would leave them, with some made to fail each gate."""
    th = rng.uniform(-0.2, 0.2, 3)
    Rx = np.array([[1, 0, 0], [0, np.cos(th[0]), -np.sin(th[0])], [0, np.sin(th[0]), np.cos(th[0])]])
    Ry = np.array([[np.cos(th[1]), 0, np.sin(th[1])], [0, 1, 0], [-np.sin(th[1]), 0, np.cos(th[1])]])
    Rz = np.array([[np.cos(th[2]), -np.sin(th[2]), 0], [np.sin(th[2]), np.cos(th[2]), 0], [0, 0, 1]])
    R = (Rx @ Ry @ Rz).astype(np.float32)
    t = rng.uniform(-1, 1, 3).astype(np.float32)
    Ow = (-R.T.astype(np.float64) @ t.astype(np.float64)).astype(np.float32)
    src = rng.integers(0, len(kps), nmp)
    z = rng.uniform(1.0, 30.0, nmp)
    xc = (kps["x"][src] + rng.normal(0, 1.0, nmp) - cam["cx"]) / cam["fx"] * z
    yc = (kps["y"][src] + rng.normal(0, 1.0, nmp) - cam["cy"]) / cam["fy"] * z
    Pc = np.stack([xc, yc, z], 1)
    scatter = rng.random(nmp) < 0.3
    Pc[scatter] = rng.uniform(-20, 20, (int(scatter.sum()), 3))
    Pw = ((Pc - t.astype(np.float64)) @ R.astype(np.float64)).astype(np.float32)        # R^T (Pc - t)
    PO = Pw.astype(np.float64) - Ow.astype(np.float64)
    dist = np.linalg.norm(PO, axis=1)
    nrm = PO / np.maximum(dist[:, None], 1e-9) + rng.normal(0, 0.35, (nmp, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None]
    flip = rng.random(nmp) < 0.1
    nrm[flip] *= -1                                                                     # viewing-angle gate
    lvl_ref = rng.integers(0, nlevels, nmp)
    dmax = dist * (1.2 ** lvl_ref) * rng.uniform(0.9, 1.3, nmp)
    dmin = dmax / (1.2 ** (nlevels - 1))
    bad = rng.random(nmp) < 0.1
    dmax[bad] = dist[bad] * 0.7                                                         # distance gate
    d = _flip(rng, desc[src].copy(), 0, 35)
    pts = dict(pos=Pw, normal=nrm.astype(np.float32), min_dist=dmin.astype(np.float32), max_dist=dmax.astype(np.float32), desc=d,
               candidate=(rng.random(nmp) < 0.9).astype(np.uint8))
    return dict(Rcw=R, tcw=t, Ow=Ow), pts


@pytest.mark.parametrize("stereo,far,th,seed", [(False, False, 1.0, 91), (True, False, 3.0, 92), (True, True, 5.0, 93), (False, False, 15.0, 94)])
def test_search_local_points(pkg, orc, oracle, stereo, far, th, seed):
    """Tracking::SearchLocalPoints: isInFrustum on the device (mTrack* fields bit-exact, incl. PredictScale through the
    restated glibc logf) feeding SearchByProjection without leaving the device."""
    rng = np.random.default_rng(seed)
    rows, cols, n, nmp = 480, 752, 1500, 5000
    cam = dict(fx=435.2047, fy=435.2047, cx=367.4517, cy=252.2008, mbf=47.90639)
    kps, desc, _, _ = _frame(pkg, rng, n, rows, cols)
    pose, pts = _scene(rng, kps, desc, rows, cols, cam, nmp)
    uR = np.where(rng.random(n) < 0.6, kps["x"] - rng.uniform(0.5, 40, n), -1.0).astype(np.float32) if stereo else None
    bounds = (0.0, float(cols), 0.0, float(rows))
    log_sf = float(np.log(np.float32(1.2)).astype(np.float32))
    pre = np.where(rng.random(n) < 0.1, 7, -1).astype(np.int32)                        # keypoints that already hold a map point
    nin_o, f_o = orc.is_in_frustum(oracle, pose, cam, bounds, log_sf, 8, pts, 0.5)
    mp = dict(desc=pts["desc"], projx=f_o["projx"], projy=f_o["projy"], projxr=f_o["projxr"], level=f_o["level"],
              viewcos=f_o["viewcos"], depth=f_o["depth"], inview=f_o["inview"])
    g = orc.Grid(oracle, kps, *bounds)
    nm_o, as_o = orc.search_by_projection(g, desc, uR, SF, mp, th, 0.8, far, 12.0, pre)
    nm_g, as_g, f_g = pkg.SearchLocalPoints(kps, desc, uR, bounds, SF, pose, cam, log_sf, pts, 0.5, th, 0.8, far, 12.0, pre)
    assert 500 < nin_o < nmp, nin_o
    for k in ("inview", "level", "projx", "projy", "projxr", "depth", "viewcos"):
        assert np.array_equal(f_g[k], f_o[k]), f"{k}: {np.sum(f_g[k] != f_o[k])} differ"
    assert len(np.unique(f_o["level"][f_o["inview"] != 0])) >= 6
    assert nm_o > 100, nm_o
    assert nm_g == nm_o and np.array_equal(as_g, as_o)


def test_predict_scale_boundaries(pkg, orc, oracle):
    """PredictScale on knife edges: max_dist / dist sitting exactly on, one ulp below and one ulp above powers of
    the scale factor (the ceil of logf(ratio) / logf(1.2f) flips there) - device logf must equal glibc's bit for bit."""
    log_sf = float(np.log(np.float32(1.2)).astype(np.float32))
    cam = dict(fx=400.0, fy=400.0, cx=320.0, cy=240.0, mbf=40.0)
    pose = dict(Rcw=np.eye(3, dtype=np.float32), tcw=np.zeros(3, np.float32), Ow=np.zeros(3, np.float32))
    base = np.float32(1.2) ** np.arange(0, 9).astype(np.float32)
    ratios = []
    for b in base:
        x = np.float32(b)
        for _ in range(40):
            x = np.nextafter(x, np.float32(0))
        for _ in range(80):
            ratios.append(x); x = np.nextafter(x, np.float32(np.inf))
    rng = np.random.default_rng(3)
    ratios = np.concatenate([np.array(ratios, np.float32), rng.uniform(0.3, 6.0, 20000).astype(np.float32)])
    m = len(ratios)
    dist = rng.uniform(2.0, 20.0, m).astype(np.float32)
    pos = np.zeros((m, 3), np.float32); pos[:, 2] = dist                              # on the optical axis: |P| = dist exactly
    pts = dict(pos=pos, normal=np.tile(np.array([0, 0, 1], np.float32), (m, 1)), min_dist=np.zeros(m, np.float32),
               max_dist=(ratios * dist).astype(np.float32), desc=np.zeros((m, 32), np.uint8), candidate=np.ones(m, np.uint8))
    bounds = (0.0, 640.0, 0.0, 480.0)
    _, f_o = orc.is_in_frustum(oracle, pose, cam, bounds, log_sf, 8, pts, 0.5)
    empty = np.zeros(0, pkg.KP_DTYPE)
    _, _, f_g = pkg.SearchLocalPoints(empty, np.zeros((0, 32), np.uint8), None, bounds, SF, pose, cam, log_sf, pts)
    assert f_o["inview"].sum() > m // 2
    assert np.array_equal(f_g["level"], f_o["level"]) and np.array_equal(f_g["inview"], f_o["inview"])
    assert np.array_equal(f_g["depth"], f_o["depth"])


@pytest.mark.parametrize("th,far,seed,preheld", [(1.0, False, 101, 0.0), (3.0, False, 102, 0.1), (5.0, True, 103, 0.1), (15.0, False, 104, 0.3)])
def test_search_by_projection_rig(pkg, orc, oracle, th, far, seed, preheld):
    """C3: SearchByProjection(Frame&, vector<MapPoint*>&, ...) on a two-camera rig (Nleft != -1): left and right keypoint
    sets, stereo partners overwritten, a failed left ratio test skipping the right side, right radius without th."""
    rng = np.random.default_rng(seed)
    rows = cols = 512
    nL, nR, nmp = 1000, 950, 5000
    kL, dL, pool, _ = _frame(pkg, rng, nL, rows, cols)
    kR, dR, _, _ = _frame(pkg, rng, nR, rows, cols, pool=pool)
    desc = np.concatenate([dL, dR])
    # stereo partners (ComputeStereoFishEyeMatches): a random partial matching
    l2r = np.full(nL, -1, np.int32); r2l = np.full(nR, -1, np.int32)
    pairs = min(nL, nR) // 3
    li = rng.choice(nL, pairs, replace=False); ri = rng.choice(nR, pairs, replace=False)
    l2r[li] = ri; r2l[ri] = li
    srcL = rng.integers(0, nL, nmp); srcR = rng.integers(0, nR, nmp)
    mp = dict(desc=_flip(rng, np.where((rng.random(nmp) < 0.5)[:, None], dL[srcL], dR[srcR]), 0, 30),
              depth=rng.uniform(0.5, 30.0, nmp).astype(np.float32))
    for s, k, src in (("L", kL, srcL), ("R", kR, srcR)):
        near = rng.random(nmp) < 0.7
        mp["projx" + s] = np.where(near, k["x"][src] + rng.normal(0, 2, nmp), rng.uniform(0, cols, nmp)).astype(np.float32)
        mp["projy" + s] = np.where(near, k["y"][src] + rng.normal(0, 2, nmp), rng.uniform(0, rows, nmp)).astype(np.float32)
        mp["level" + s] = np.clip(k["octave"][src] + rng.integers(0, 2, nmp), 0, 7).astype(np.int32)
        mp["viewcos" + s] = np.where(rng.random(nmp) < 0.5, 0.9, 0.999).astype(np.float32)
        mp["inview" + s] = (rng.random(nmp) < 0.6).astype(np.uint8)
    mp["levelR"][rng.random(nmp) < 0.05] = -1                                     # mnTrackScaleLevelR == -1, :172
    pre = np.where(rng.random(nL + nR) < preheld, 3, -1).astype(np.int32)
    bounds = (0.0, float(cols), 0.0, float(rows))
    gL, gR = orc.Grid(oracle, kL, *bounds), orc.Grid(oracle, kR, *bounds)
    nm_o, as_o = orc.search_by_projection_rig(gL, gR, desc, l2r, r2l, SF, mp, th, 0.8, far, 12.0, pre)
    nm_g, as_g = pkg.SearchByProjectionRig(kL, kR, desc, l2r, r2l, bounds, SF, mp, th, 0.8, far, 12.0, pre)
    assert nm_o > 300, nm_o
    assert nm_g == nm_o
    assert np.array_equal(as_g, as_o), f"{np.sum(as_g != as_o)} assignments differ"
    # a pre-held keypoint can only change through a partner overwrite
    changed = (pre >= 0) & (as_o != pre)
    partners = np.concatenate([r2l >= 0, l2r >= 0]) if False else np.concatenate([l2r >= 0, r2l >= 0])
    assert not np.any(changed & ~partners)

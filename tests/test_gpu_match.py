"""GPU parity of the matching path (stereo row search + SAD, brute-force 2-NN, SearchByProjection)
against the CPU oracle.  Integer results bit-exact; mvuRight/mvDepth bit-exact floats."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _stereo_case(pkg, orc, oracle, cfg, seed, nframes=1, mb=None):
    c = pkg.synth.CONFIGS[cfg]
    rows, cols, nf = c["rows"], c["cols"], c["nfeatures"]
    mbf = c["mbf"]; mb = mbf / c["fx"] if mb is None else mb
    pairs = [pkg.synth.synth_stereo_pair(rows, cols, seed + i) for i in range(nframes)]
    L = np.stack([p[0] for p in pairs]); R = np.stack([p[1] for p in pairs])
    XL, XR = pkg.ORBextractor(nf), pkg.ORBextractor(nf)
    _, nL, kL, dL = XL.extract_batch(L, (0, 0))
    _, nR, kR, dR = XR.extract_batch(R, (0, 0))
    uR, depth = pkg.ComputeStereoMatches(XL, XR, mb, mbf, frame=None)
    EL, ER = orc.Extractor(nf, oracle=oracle), orc.Extractor(nf, oracle=oracle)
    matched = 0
    for f in range(nframes):
        _, kl, dl = EL(L[f], (0, 0)); _, kr, dr = ER(R[f], (0, 0))
        assert np.array_equal(kL[f, :nL[f]], kl) and np.array_equal(kR[f, :nR[f]], kr)
        u_o, d_o, _, _ = orc.stereo_match(EL, ER, kl, dl, kr, dr, mb, mbf)
        assert np.array_equal(uR[f, :nL[f]], u_o), f"mvuRight frame {f}: {np.sum(uR[f, :nL[f]] != u_o)} differ"
        assert np.array_equal(depth[f, :nL[f]], d_o), f"mvDepth frame {f}"
        matched += int((u_o >= 0).sum())
    XL.close(); XR.close()
    return matched


def test_stereo_euroc(pkg, orc, oracle):
    assert _stereo_case(pkg, orc, oracle, "C2", 11, nframes=3) > 300     # the synthetic pairs do match


def test_stereo_kitti(pkg, orc, oracle):
    assert _stereo_case(pkg, orc, oracle, "C4", 21, nframes=2) > 300


def test_stereo_unbounded_disparity(pkg, orc, oracle):
    """mb = 0: maxD = mbf / mb = +inf.  This is what the reference's rectified-stereo constructor effectively runs with - it calls
    ComputeStereoMatches (src/Frame.cc:137) before it assigns mb (:163); tests/test_oracle_ref_frame.py::test_stereo_constructor_end_to_end."""
    assert _stereo_case(pkg, orc, oracle, "C4", 23, nframes=2, mb=0.0) > 300


@pytest.mark.parametrize("nframes", [9, 11])
def test_stereo_batch_odd_sizes_stay_resident(pkg, orc, oracle, nframes):
    """Batch sizes that do not divide into equal chunks (9 = 3+3+3, 11 = 3+3+3+2): every frame must sit in its own
    device slot after orbx_extract_batch, for orbx_stereo_match_batch and the pyramid getter alike."""
    assert _stereo_case(pkg, orc, oracle, "C2", 40 + nframes, nframes=nframes) > 100 * nframes


def test_stereo_after_handle_growth(pkg, orc, oracle):
    """A handle first used for ONE frame must serve a later 6-pair batch with all pairs resident (the slots grow)."""
    c = pkg.synth.CONFIGS["C2"]
    rows, cols, nf = c["rows"], c["cols"], c["nfeatures"]
    mbf = c["mbf"]; mb = mbf / c["fx"]
    XL, XR = pkg.ORBextractor(nf), pkg.ORBextractor(nf)
    l0, r0 = pkg.synth.synth_stereo_pair(rows, cols, 70)
    XL(l0); XR(r0)
    pairs = [pkg.synth.synth_stereo_pair(rows, cols, 71 + i) for i in range(6)]
    L = np.stack([p[0] for p in pairs]); R = np.stack([p[1] for p in pairs])
    _, nL, kL, dL = XL.extract_batch(L, (0, 0))
    _, nR, kR, dR = XR.extract_batch(R, (0, 0))
    uR, depth = pkg.ComputeStereoMatches(XL, XR, mb, mbf, frame=None)
    EL, ER = orc.Extractor(nf, oracle=oracle), orc.Extractor(nf, oracle=oracle)
    for f in (0, 5):
        _, kl, dl = EL(L[f], (0, 0)); _, kr, dr = ER(R[f], (0, 0))
        u_o, d_o, _, _ = orc.stereo_match(EL, ER, kl, dl, kr, dr, mb, mbf)
        assert np.array_equal(uR[f, :nL[f]], u_o) and np.array_equal(depth[f, :nL[f]], d_o)
        assert np.array_equal(XL.pyramid_level(3, frame=f), EL.level_plane(3)[19:-19, 19:-19])
    XL.close(); XR.close()


def test_stereo_single_frame_api(pkg, orc, oracle):
    c = pkg.synth.CONFIGS["C2"]
    l, r = pkg.synth.synth_stereo_pair(c["rows"], c["cols"], 5)
    XL, XR = pkg.ORBextractor(1200), pkg.ORBextractor(1200)
    _, kl, dl = XL(l); _, kr, dr = XR(r)
    mb = c["mbf"] / c["fx"]
    uR, depth = pkg.ComputeStereoMatches(XL, XR, mb, c["mbf"], frame=0)
    EL, ER = orc.Extractor(1200, oracle=oracle), orc.Extractor(1200, oracle=oracle)
    _, ko, do = EL(l); _, kro, dro = ER(r)
    u_o, d_o, _, _ = orc.stereo_match(EL, ER, ko, do, kro, dro, mb, c["mbf"])
    assert np.array_equal(uR[:len(kl)], u_o) and np.array_equal(depth[:len(kl)], d_o)
    XL.close(); XR.close()


@pytest.mark.parametrize("cfg,seed,strided", [("C2", 5, False), ("C4", 9, True)])
def test_extract_stereo_one_call(pkg, orc, oracle, cfg, seed, strided):
    """orbx_extract_stereo (both eyes + ComputeStereoMatches in one call) == the oracle, and the handles are left in
    the same state as after separate calls (a following ComputeStereoMatches gives the same answer).  Repeated on the
    same handles with another pair; strided (non-continuous) input rows."""
    c = pkg.synth.CONFIGS[cfg]
    nf = c["nfeatures"]; mbf = c["mbf"]; mb = mbf / c["fx"]
    XL, XR = pkg.ORBextractor(nf), pkg.ORBextractor(nf)
    EL, ER = orc.Extractor(nf, oracle=oracle), orc.Extractor(nf, oracle=oracle)
    for rep in range(2):
        l, r = pkg.synth.synth_stereo_pair(c["rows"], c["cols"], seed + rep)
        if strided:
            big = np.zeros((2, c["rows"], c["cols"] + 24), np.uint8)
            big[0, :, :c["cols"]] = l; big[1, :, :c["cols"]] = r
            lv, rv = big[0, :, :c["cols"]], big[1, :, :c["cols"]]
        else:
            lv, rv = l, r
        kl, dl, kr, dr, uR, depth = pkg.ExtractStereo(XL, XR, lv, rv, mb, mbf)
        _, ko, do = EL(l); _, kro, dro = ER(r)
        assert np.array_equal(kl, ko) and np.array_equal(dl, do) and np.array_equal(kr, kro) and np.array_equal(dr, dro)
        u_o, d_o, _, _ = orc.stereo_match(EL, ER, ko, do, kro, dro, mb, mbf)
        assert np.array_equal(uR, u_o) and np.array_equal(depth, d_o)
        assert (u_o >= 0).sum() > 100
        u2, d2 = pkg.ComputeStereoMatches(XL, XR, mb, mbf, frame=0)
        assert np.array_equal(u2[:len(kl)], u_o) and np.array_equal(d2[:len(kl)], d_o)
    XL.close(); XR.close()


def test_extract_stereo_flat_frames(pkg, orc, oracle):
    """No keypoints in one or both eyes (src/Frame.cc:145 returns before ComputeStereoMatches): empty results, no
    error, and the handles still work afterwards."""
    c = pkg.synth.CONFIGS["C2"]
    mbf = c["mbf"]; mb = mbf / c["fx"]
    flat = np.full((c["rows"], c["cols"]), 77, np.uint8)
    l, r = pkg.synth.synth_stereo_pair(c["rows"], c["cols"], 15)
    XL, XR = pkg.ORBextractor(600), pkg.ORBextractor(600)
    kl, dl, kr, dr, uR, depth = pkg.ExtractStereo(XL, XR, flat, flat, mb, mbf)
    assert len(kl) == 0 and len(kr) == 0 and len(uR) == 0 and dl.shape == (0, 32)
    kl, dl, kr, dr, uR, depth = pkg.ExtractStereo(XL, XR, l, flat, mb, mbf)        # nothing to match against
    assert len(kl) > 100 and len(kr) == 0 and np.all(uR == -1) and np.all(depth == -1)
    kl, dl, kr, dr, uR, depth = pkg.ExtractStereo(XL, XR, l, r, mb, mbf)
    EL, ER = orc.Extractor(600, oracle=oracle), orc.Extractor(600, oracle=oracle)
    _, ko, do = EL(l); _, kro, dro = ER(r)
    u_o, d_o, _, _ = orc.stereo_match(EL, ER, ko, do, kro, dro, mb, mbf)
    assert np.array_equal(kl, ko) and np.array_equal(uR, u_o) and np.array_equal(depth, d_o)
    XL.close(); XR.close()


def test_extract_stereo_errors(pkg):
    XL, XR = pkg.ORBextractor(500), pkg.ORBextractor(500)
    im = pkg.synth.synth_frame(480, 752, 3)
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ExtractStereo(XL, XL, im, im, 0.11, 47.9)            # the same handle twice
    assert ei.value.code == pkg.binding.ORBX_ERR_ARG
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ExtractStereo(XL, XR, im, im, -0.1, 47.9)            # mb must be >= 0 (0 = no upper disparity bound)
    assert ei.value.code == pkg.binding.ORBX_ERR_ARG
    XL.close(); XR.close()


def test_stereo_state_errors(pkg):
    XL, XR = pkg.ORBextractor(1200), pkg.ORBextractor(1200)
    XL._last_shape = (1, 480, 752)
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ComputeStereoMatches(XL, XR, 0.11, 47.9)
    assert ei.value.code == pkg.binding.ORBX_ERR_STATE
    XL.close(); XR.close()


@pytest.mark.parametrize("nq,nt", [(1000, 1000), (517, 733), (1, 1), (3, 2), (5, 0), (1300, 129)])
def test_knn2(pkg, oracle, nq, nt):
    rng = np.random.default_rng(nq * 7 + nt)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    if nt > 4:      # forced duplicates / near-duplicates: tie-break by ascending train index
        t[nt // 2] = t[1]; t[nt - 1] = t[1]
        q[0] = t[1]
        q[nq // 2] = t[1] ^ np.uint8(1)
    gi, gd = pkg.knn2_hamming(q, t)
    oi, od = oracle.knn2(q, t) if nt > 0 else (np.full((nq, 2), -1, np.int32), np.full((nq, 2), -1, np.int32))
    assert np.array_equal(gi, oi) and np.array_equal(gd, od)


def test_descriptor_distance(pkg, oracle):
    rng = np.random.default_rng(1)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert pkg.DescriptorDistance(a, b) == oracle.hamming(a, b) == int(np.unpackbits(a ^ b).sum())


@pytest.mark.parametrize("th,stereo,far", [(1.0, False, False), (3.0, False, False), (3.0, True, False), (1.0, True, True)])
def test_search_by_projection(pkg, orc, oracle, th, stereo, far):
    """C3: 512x512 fisheye-sized frame vs 5000 synthetic map points."""
    rows, cols = 512, 512
    img = pkg.synth.synth_frame(rows, cols, 31)
    X = pkg.ORBextractor(1000)
    _, kps, desc = X(img, (0, 0))
    sf = X.GetScaleFactors()
    mp = pkg.synth.synth_map_points(desc, np.stack([kps["x"], kps["y"]], 1), kps["octave"], 5000, 77, rows, cols)
    rng = np.random.default_rng(5)
    uR = None
    if stereo:
        uR = np.where(rng.random(len(kps)) < 0.6, kps["x"] - rng.uniform(0, 30, len(kps)), -1).astype(np.float32)
    pre = np.full(len(kps), -1, np.int32)
    pre[rng.choice(len(kps), 50, replace=False)] = 9999          # keypoints that already hold a map point
    g = orc.Grid(oracle, kps, 0.0, float(cols), 0.0, float(rows))
    nm_o, as_o = orc.search_by_projection(g, desc, uR, sf, mp, th, 0.8, far, 30.0, pre)
    nm_g, as_g = pkg.SearchByProjection(kps, desc, uR, (0.0, float(cols), 0.0, float(rows)), sf, mp, th, 0.8, far, 30.0, pre)
    assert nm_o > 100
    assert nm_g == nm_o
    assert np.array_equal(as_g, as_o), f"{np.sum(as_g != as_o)} assignments differ"
    X.close()


def test_search_by_projection_long_candidate_lists(pkg, orc, oracle):
    """th = 15 right after relocalisation on a dense frame: a map point's window holds far more than the 256 keypoints the
    candidate lists start with; the reference has no limit, the lists are regrown and the pass repeated."""
    rng = np.random.default_rng(17)
    rows, cols, n = 480, 752, 6000
    kps = np.zeros(n, pkg.KP_DTYPE)
    kps["x"] = rng.uniform(250, 500, n).astype(np.float32); kps["y"] = rng.uniform(150, 330, n).astype(np.float32)   # a crowded region
    kps["octave"] = rng.choice(8, n, p=[.3, .22, .16, .12, .08, .06, .04, .02]); kps["angle"] = 0; kps["class_id"] = -1
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    sf = np.cumprod(np.concatenate([[1.0], np.full(7, 1.2)])).astype(np.float32)
    m = 800
    src = rng.integers(0, n, m)
    mp = dict(desc=desc[src] ^ (rng.random((m, 32)) < 0.03).astype(np.uint8),
              projx=(kps["x"][src] + rng.normal(0, 4, m)).astype(np.float32), projy=(kps["y"][src] + rng.normal(0, 4, m)).astype(np.float32),
              projxr=np.zeros(m, np.float32), level=np.clip(kps["octave"][src] + 1, 0, 7).astype(np.int32),
              viewcos=np.full(m, 0.9, np.float32), depth=np.full(m, 5.0, np.float32), inview=np.ones(m, np.uint8))
    b = (0.0, float(cols), 0.0, float(rows))
    g = orc.Grid(oracle, kps, *b)
    longest = max(len(g.query(float(mp["projx"][i]), float(mp["projy"][i]), 4.0 * 15.0 * float(sf[mp["level"][i]]), int(mp["level"][i]) - 1, int(mp["level"][i])))
                  for i in range(0, m, 16))
    assert longest > 600                                   # well past the initial capacity
    for _ in range(2):                                     # second call: the grown lists are kept
        nm_o, as_o = orc.search_by_projection(g, desc, None, sf, mp, 15.0, 0.8, False, 0.0, None)
        nm_g, as_g = pkg.SearchByProjection(kps, desc, None, b, sf, mp, 15.0, 0.8, False, 0.0, None)
        assert nm_o > 100 and nm_g == nm_o and np.array_equal(as_g, as_o)


@pytest.mark.parametrize("seed,maxn", [(1, 12), (2, 40), (3, 300)])
def test_distinctive_descriptors(pkg, orc, oracle, seed, maxn):
    """MapPoint::ComputeDistinctiveDescriptors batched: groups of 0..maxn observation descriptors (ties are frequent:
    descriptors of a group are a few flips away from one another), best row and its median equal to the oracle's."""
    rng = np.random.default_rng(seed)
    npts = 700 if maxn < 100 else 60
    sizes = rng.integers(0, maxn + 1, npts)
    sizes[:4] = [0, 1, 2, maxn]
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = rng.integers(0, 256, (npts, 32), dtype=np.uint8)
    desc = np.repeat(base, sizes, axis=0)
    bits = np.unpackbits(desc, axis=1)
    flips = rng.random(bits.shape) < rng.choice([0.0, 0.02, 0.1], (len(bits), 1))
    desc = np.packbits(bits ^ flips.astype(np.uint8), axis=1)
    bi_o, bm_o = orc.distinctive_descriptors(oracle, start, desc)
    bi_g, bm_g = pkg.ComputeDistinctiveDescriptors(start, desc)
    assert np.array_equal(bi_g, bi_o), f"{np.sum(bi_g != bi_o)} best rows differ"
    assert np.array_equal(bm_g, bm_o)
    assert bi_o[0] == -1 and bi_o[1] == 0


@pytest.mark.parametrize("k,L,levelsup,seed", [(10, 4, 2, 1), (10, 4, 4, 2), (6, 5, 4, 3), (20, 3, 1, 4)])
def test_bow_transform(pkg, orc, oracle, k, L, levelsup, seed):
    """Frame::ComputeBoW: vocabulary-tree descent per feature (word, weight, node at L - levelsup) and the BowVector /
    FeatureVector built from it, against the oracle's restatement of TemplatedVocabulary::transform."""
    rng = np.random.default_rng(seed)
    voc = pkg.synth.synth_vocabulary(k, L, seed)
    V = pkg.ORBVocabulary(k, L, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    leaves = voc["desc"][voc["is_leaf"] != 0]
    n = 1500
    desc = leaves[rng.integers(0, len(leaves), n)].copy()
    bits = np.unpackbits(desc, axis=1)
    desc = np.packbits(bits ^ (rng.random(bits.shape) < rng.choice([0.0, 0.05, 0.3], (n, 1))).astype(np.uint8), axis=1)
    desc[:5] = 0                                                    # equal distances to many children: first least wins
    w_o, wt_o, nid_o = orc.bow_transform(oracle, voc, desc, levelsup)
    w_g, wt_g, nid_g = V.transform(desc, levelsup)
    assert np.array_equal(w_g, w_o) and np.array_equal(nid_g, nid_o) and np.array_equal(wt_g, wt_o)
    assert (wt_o == 0).any() and len(np.unique(w_o)) > 200
    got = V.ComputeBoW(desc, levelsup)
    want = orc.compute_bow(oracle, voc, desc, levelsup)
    for a, b, name in zip(got, want, ("bow_id", "bow_val", "fv_node", "fv_start", "fv_idx")):
        assert np.array_equal(a, b), name
    assert abs(got[1].sum() - 1.0) < 1e-12
    V.close()


def test_vocabulary_argument_errors(pkg):
    import ctypes as C
    lib = pkg.binding.load()
    h = C.c_void_p()
    parent = np.array([0, 0, 5], np.int32); leaf = np.array([1, 1, 1], np.uint8)
    d = np.zeros((3, 32), np.uint8); w = np.ones(3, np.float64)
    assert lib.orbx_vocabulary_create(0, 10, 2, 3, parent.ctypes.data, leaf.ctypes.data, d.ctypes.data, w.ctypes.data, C.byref(h)) == pkg.binding.ORBX_ERR_ARG
    parent = np.array([0, 0, 1], np.int32)                          # node 1 has a child but is flagged as a leaf
    assert lib.orbx_vocabulary_create(0, 10, 2, 3, parent.ctypes.data, leaf.ctypes.data, d.ctypes.data, w.ctypes.data, C.byref(h)) == pkg.binding.ORBX_ERR_ARG


@pytest.mark.parametrize("ncoef,seed", [(4, 1), (5, 2)])
def test_undistort_keypoints(pkg, oracle, ncoef, seed):
    """Frame::UndistortKeyPoints on the device: positions bit-exact against the oracle (itself pinned to cv2), every
    other KeyPoint field untouched; k1 == 0 is the identity."""
    rng = np.random.default_rng(seed)
    n = 3000
    kps = np.zeros(n, pkg.KP_DTYPE)
    kps["x"] = rng.uniform(0, 752, n); kps["y"] = rng.uniform(0, 480, n)
    kps["octave"] = rng.integers(0, 8, n); kps["angle"] = rng.uniform(0, 360, n); kps["response"] = rng.integers(7, 200, n); kps["size"] = 31; kps["class_id"] = -1
    cam = (458.654, 457.296, 367.215, 248.375)                        # EuRoC cam0 (Examples/Monocular/EuRoC.yaml)
    dist = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.01][:ncoef], np.float32)
    un = pkg.UndistortKeyPoints(kps, *cam, dist)
    want = oracle.undistort_points(np.stack([kps["x"], kps["y"]], 1), *cam, dist)
    assert np.array_equal(un["x"], want[:, 0]) and np.array_equal(un["y"], want[:, 1])
    assert np.abs(un["x"] - kps["x"]).max() > 1.0
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(un[f], kps[f])
    dist0 = dist.copy(); dist0[0] = 0
    assert np.array_equal(pkg.UndistortKeyPoints(kps, *cam, dist0), kps)

"""GPU parity (through the C ABI) of the two-camera-rig branches (Frame::Nleft != -1: the KannalaBrandt stereo set-up of
config C3) that SURVEY.md §8 row 16 lists, the triangulation searches whose geometric verdict is a caller-side callback,
and Frame::ComputeStereoFishEyeMatches on device-resident descriptors — each against the oracle restatement on the same
seeded inputs.  Bar: identical match vectors and counts; identical callback call SEQUENCES."""
import numpy as np
import pytest

from test_gpu_search import SF, _feature_vector, _flip, _frame, _moved

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("direction,orient,seed,th", [("none", True, 71, 15.0), ("forward", True, 72, 7.0),
                                                     ("backward", False, 73, 7.0), ("none", True, 74, 3.0)])
def test_search_by_projection_frame_rig(pkg, orc, oracle, direction, orient, seed, th):
    rng = np.random.default_rng(seed)
    rows, cols = 512, 512
    nL, nR = 1100, 900
    lk, ld, _, _ = _frame(pkg, rng, nL, rows, cols)                  # last frame's keypoints (rows of the search)
    ckL, cdL, _ = _moved(pkg, rng, lk, ld, rows, cols)               # current left camera: the same points, moved
    ckR, cdR, _ = _moved(pkg, rng, lk[:nR], ld[:nR], rows, cols, shift=(-20.0, 1.0))
    desc = np.concatenate([cdL, cdR])
    n = nL
    last = dict(desc=_flip(rng, ld.copy(), 0, 10),
                u=(lk["x"] + 4.0 + rng.normal(0, 3, n)).astype(np.float32), v=(lk["y"] - 3.0 + rng.normal(0, 3, n)).astype(np.float32),
                ur=(lk["x"] - 20.0 + rng.normal(0, 3, n)).astype(np.float32), vr=(lk["y"] + 1.0 + rng.normal(0, 3, n)).astype(np.float32),
                invz=np.where(rng.random(n) < 0.97, rng.uniform(0.02, 0.5, n), -0.1).astype(np.float32),
                octave=lk["octave"].astype(np.int32), angle=lk["angle"],
                valid=(rng.random(n) < 0.85).astype(np.uint8), obs=(rng.random(n) < 0.9).astype(np.uint8))
    # some left windows are empty although the right one is not (the `continue` at :2541 skips both halves)
    far = rng.random(n) < 0.08
    last["u"][far] = np.where(rng.random(int(far.sum())) < 0.5, 2.0, cols - 2.0)
    last["v"][far] = rows - 1.0
    held = (rng.random(nL + nR) < 0.03).astype(np.uint8)
    fwd, bwd = direction == "forward", direction == "backward"
    bounds = (0.0, float(cols), 0.0, float(rows))
    gL, gR = orc.Grid(oracle, ckL, *bounds), orc.Grid(oracle, ckR, *bounds)
    nm_o, as_o = orc.search_by_projection_frame_rig(gL, gR, desc, SF, last, th, fwd, bwd, orient, held)
    nm_g, as_g = pkg.SearchByProjectionFrameRig(ckL, ckR, desc, bounds, SF, last, th, fwd, bwd, orient, held)
    assert nm_o > 150 and int(np.sum(as_o[nL:] >= 0)) > 50 and int(np.sum(as_o[:nL] >= 0)) > 50
    assert nm_g == nm_o
    assert np.array_equal(as_g, as_o), f"{np.sum(as_g != as_o)} assignments differ"
    # no right camera keypoints at all / no left ones
    e = np.zeros(0, pkg.KP_DTYPE)
    nm_o, as_o = orc.search_by_projection_frame_rig(gL, orc.Grid(oracle, e, *bounds), cdL, SF, last, th, fwd, bwd, orient, held[:nL])
    nm_g, as_g = pkg.SearchByProjectionFrameRig(ckL, e, cdL, bounds, SF, last, th, fwd, bwd, orient, held[:nL])
    assert nm_g == nm_o and np.array_equal(as_g, as_o)
    nm_g, as_g = pkg.SearchByProjectionFrameRig(e, ckR, cdR, bounds, SF, last, th, fwd, bwd, orient, None)
    assert nm_g == 0 and np.all(as_g == -1)             # every left window is empty -> the right half never runs


def _bow_pair(pkg, rng, n, flips=(0, 50), nodes=89, p_valid=0.7):
    rows, cols = 512, 512
    k1, d1, pool, src = _frame(pkg, rng, n, rows, cols)
    k2, d2, perm = _moved(pkg, rng, k1, d1, rows, cols, flips=flips)
    node1 = (src % nodes).astype(np.int32); node2 = node1[perm]
    node2 = np.where(rng.random(n) < 0.05, rng.integers(0, nodes, n), node2).astype(np.int32)
    kf1 = dict(kps=k1, desc=d1)
    kf2 = dict(kps=k2, desc=d2)
    kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = _feature_vector(rng, node1, 0.02)
    kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = _feature_vector(rng, node2, 0.02)
    return kf1, kf2


@pytest.mark.parametrize("nnratio,orient,seed,nleft_frac", [(0.7, True, 81, 0.55), (0.9, False, 82, 0.5), (0.75, True, 83, 0.0),
                                                           (0.7, True, 84, 1.0), (1.3, True, 85, 0.4)])
def test_search_by_bow_rig(pkg, orc, oracle, nnratio, orient, seed, nleft_frac):
    rng = np.random.default_rng(seed)
    n = 1600
    kf1, kf2 = _bow_pair(pkg, rng, n)
    kf1["valid"] = (rng.random(n) < 0.7).astype(np.uint8)
    nleft2 = int(n * nleft_frac)
    nm_o, m_o = orc.search_by_bow_rig(oracle, kf1, kf2, nleft2, nnratio, orient)
    nm_g, m_g = pkg.SearchByBoWRig(kf1, kf2, nleft2, nnratio, orient)
    if 0 < nleft2 < n:
        assert nm_o > 40 and int(np.sum(m_o[nleft2:] >= 0)) > 10 and int(np.sum(m_o[:nleft2] >= 0)) > 10
    assert nm_g == nm_o
    assert np.array_equal(m_g, m_o), f"{np.sum(m_g != m_o)} matches differ"
    if nleft2 == n:      # no right-camera rows: must equal the Nleft == -1 branch
        nm_p, m_p = pkg.SearchByBoW(0, dict(kf1), dict(kf2, valid=None), nnratio, orient)
        assert nm_p == nm_g and np.array_equal(m_p, m_g)


def _verdict_fn(seed, p):
    """deterministic pseudo-random stand-in for the camera model's test; records the call sequence"""
    calls = []

    def v(i1, i2):
        calls.append((i1, i2))
        h = (i1 * 2654435761 + i2 * 40503 + seed * 97) & 0xFFFFFFFF
        h ^= h >> 13; h = (h * 1274126177) & 0xFFFFFFFF; h ^= h >> 16
        return (h % 1000) < int(p * 1000)
    return v, calls


@pytest.mark.parametrize("overload,only_stereo,coarse,orient,p,seed", [(0, False, False, True, 0.6, 91), (1, False, False, True, 0.6, 92),
                                                                      (2, False, False, True, 0.5, 93), (0, False, True, True, 0.5, 94),
                                                                      (1, True, False, True, 0.5, 95), (2, True, True, False, 0.3, 96),
                                                                      (0, False, False, False, 1.0, 97), (2, False, False, True, 0.0, 98)])
def test_search_for_triangulation_callback(pkg, orc, oracle, overload, only_stereo, coarse, orient, p, seed):
    rng = np.random.default_rng(seed)
    n = 1500
    kf1, kf2 = _bow_pair(pkg, rng, n, flips=(0, 45), nodes=97)
    kf1["has_mp"] = (rng.random(n) < 0.35).astype(np.uint8)
    kf2["has_mp"] = (rng.random(n) < 0.35).astype(np.uint8)
    vo, calls_o = _verdict_fn(seed, p)
    vg, calls_g = _verdict_fn(seed, p)
    nm_o, m_o = orc.search_for_triangulation_cb(oracle, overload, kf1, kf2, vo, only_stereo, coarse, orient)
    nm_g, m_g = pkg.SearchForTriangulationCb(overload, kf1, kf2, vg, only_stereo, coarse, orient)
    assert nm_g == nm_o and np.array_equal(m_g, m_o), f"{np.sum(m_g != m_o)} matches differ"
    assert calls_g == calls_o                             # the camera model is asked for the same pairs in the same order
    if overload != 2 and only_stereo:
        assert nm_o == 0 and not calls_o                  # bStereo is false with a rig: nothing to match (:1194)
    elif overload != 2 and coarse:
        assert nm_o > 40 and len(calls_o) >= nm_o         # `A || bCoarse`: the camera model is still asked, its answer ignored
    elif p > 0:
        assert nm_o > 40 and len(calls_o) > nm_o
    else:
        assert nm_o == 0 and len(calls_o) > 100


@pytest.mark.parametrize("lap,seed", [((0, 511), 5), ((128, 383), 6), ((300, 310), 7), ((0, 0), 8)])
def test_stereo_fisheye_resident(pkg, orc, oracle, lap, seed):
    """C3: two extractions with vLappingArea, then the kNN + ratio test on the descriptors they left on the device."""
    c = pkg.synth.CONFIGS["C3"]
    rows, cols, nf = c["rows"], c["cols"], c["nfeatures"]
    left, right = pkg.synth.synth_stereo_pair(rows, cols, seed, max_disp=30)
    XL, XR = pkg.ORBextractor(nf), pkg.ORBextractor(nf)
    monoL, kL, dL = XL(left, lap)
    monoR, kR, dR = XR(right, lap)
    good_g, l2r_g, r2l_g, bd_g = pkg.ComputeStereoFishEyeMatches(XL, XR, len(kL), monoL, len(kR), monoR)
    good_o, l2r_o, r2l_o, bd_o = orc.stereo_fisheye_match(oracle, dL, monoL, dR, monoR)
    assert good_g == good_o == int(np.sum(l2r_o >= 0))
    assert np.array_equal(l2r_g, l2r_o) and np.array_equal(r2l_g, r2l_o) and np.array_equal(bd_g, bd_o)
    if lap[1] - lap[0] > 100:
        assert good_o > 50
    # the same result through the host-array kNN entry point + the caller-side ratio test
    if monoL < len(kL) and monoR < len(kR):
        idx, dist = pkg.knn2_hamming(dL[monoL:], dR[monoR:])
        ok = (idx[:, 1] >= 0) & (dist[:, 0].astype(np.float64) < dist[:, 1].astype(np.float64) * 0.7)
        exp = np.full(len(kL), -1, np.int32); exp[monoL:][ok] = idx[ok, 0] + monoR
        assert np.array_equal(exp, l2r_g)
    XL.close(); XR.close()


def test_stereo_fisheye_batch_slot_and_errors(pkg, orc, oracle):
    c = pkg.synth.CONFIGS["C3"]
    rows, cols, nf = c["rows"], c["cols"], c["nfeatures"]
    pairs = [pkg.synth.synth_stereo_pair(rows, cols, 20 + i, max_disp=30) for i in range(3)]
    L = np.stack([p[0] for p in pairs]); R = np.stack([p[1] for p in pairs])
    XL, XR = pkg.ORBextractor(nf), pkg.ORBextractor(nf)
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ComputeStereoFishEyeMatches(XL, XR, 10, 0, 10, 0)
    assert ei.value.code == pkg.binding.ORBX_ERR_STATE
    monoL, nL, kL, dL = XL.extract_batch(L, (100, 400))
    monoR, nR, kR, dR = XR.extract_batch(R, (100, 400))
    for f in (2, 0):
        good_g, l2r_g, r2l_g, _ = pkg.ComputeStereoFishEyeMatches(XL, XR, nL[f], monoL[f], nR[f], monoR[f], frame=f)
        good_o, l2r_o, r2l_o, _ = orc.stereo_fisheye_match(oracle, dL[f, :nL[f]], monoL[f], dR[f, :nR[f]], monoR[f])
        assert good_g == good_o > 30 and np.array_equal(l2r_g, l2r_o) and np.array_equal(r2l_g, r2l_o)
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ComputeStereoFishEyeMatches(XL, XR, nL[0], nL[0] + 1, nR[0], 0)
    assert ei.value.code == pkg.binding.ORBX_ERR_ARG
    XL.close(); XR.close()

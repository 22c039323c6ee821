"""CPU checks of the oracle's guided-search restatements (oracle/orc_match.c) on cases whose answer is
known by construction, and of the host-only C-ABI entry orbx_three_maxima against the oracle and a
literal Python transcription of ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2863-2905)."""
import numpy as np

SF = np.cumprod(np.concatenate([[1.0], np.full(7, 1.2)])).astype(np.float32)


def _three_maxima_py(h):
    m1 = m2 = m3 = 0
    i1 = i2 = i3 = -1
    for i, s in enumerate(h):
        if s > m1:
            m3, m2, m1 = m2, m1, s
            i3, i2, i1 = i2, i1, i
        elif s > m2:
            m3, m2 = m2, s
            i3, i2 = i2, i
        elif s > m3:
            m3, i3 = s, i
    if m2 < np.float32(0.1) * np.float32(m1):
        i2 = i3 = -1
    elif m3 < np.float32(0.1) * np.float32(m1):
        i3 = -1
    return i1, i2, i3


def test_three_maxima(pkg, orc, oracle):
    rng = np.random.default_rng(0)
    cases = [np.zeros(30, np.int32), np.arange(30, dtype=np.int32), np.full(30, 7, np.int32)]
    for _ in range(300):
        h = rng.integers(0, rng.choice([2, 5, 50, 500]), 30).astype(np.int32)
        if rng.random() < 0.3:
            h[rng.integers(0, 30)] = 1000
        cases.append(h)
    for h in cases:
        want = _three_maxima_py(h.tolist())
        assert orc.three_maxima(oracle, h) == want
        assert pkg.ComputeThreeMaxima(h) == want


def _unique_frame(orc, rng, n, rows=480, cols=752):
    kps = np.zeros(n, orc.KP_DTYPE)
    w = cols // 8 - 2                                         # stay clear of the last grid column/row:
    cells = rng.choice((rows // 8 - 2) * w, n, replace=False)  # PosInGrid rounds, so x > 746 falls outside the grid
    kps["x"] = (cells % w) * 8 + 4
    kps["y"] = (cells // w) * 8 + 4
    kps["octave"] = 0
    kps["angle"] = 10.0
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    return kps, desc


def test_oracle_initialization_identity(orc, oracle):
    rng = np.random.default_rng(1)
    kps, desc = _unique_frame(orc, rng, 400)
    g = orc.Grid(oracle, kps, 0.0, 752.0, 0.0, 480.0)
    prev = np.stack([kps["x"], kps["y"]], 1)
    nm, m12, pm = orc.search_for_initialization(oracle, kps, desc, g, desc, prev, 100, 0.9, True)
    assert nm == 400 and np.array_equal(m12, np.arange(400))
    assert np.array_equal(pm, prev.astype(np.float32))
    kps["octave"][::2] = 1                                    # only level-0 keypoints take part (:825)
    g = orc.Grid(oracle, kps, 0.0, 752.0, 0.0, 480.0)
    nm, m12, _ = orc.search_for_initialization(oracle, kps, desc, g, desc, prev, 100, 0.9, True)
    assert nm == 200 and np.all(m12[::2] == -1) and np.array_equal(m12[1::2], np.arange(400)[1::2])


def test_oracle_projection_frame_identity_and_rotation_filter(orc, oracle):
    rng = np.random.default_rng(2)
    n = 300
    kps, desc = _unique_frame(orc, rng, n)
    g = orc.Grid(oracle, kps, 0.0, 752.0, 0.0, 480.0)
    last = dict(desc=desc, u=kps["x"], v=kps["y"], invz=np.full(n, 0.1, np.float32), octave=np.zeros(n, np.int32),
                angle=kps["angle"].copy(), valid=np.ones(n, np.uint8), obs=np.ones(n, np.uint8))
    nm, a = orc.search_by_projection_frame(g, desc, None, SF, last, 7.0, 0.0, False, False, True)
    assert nm == n and np.array_equal(a, np.arange(n))
    # 20 matches rotated by 90 degrees fall in a minority bin (< 10 % of the main one) and are dropped
    last["angle"][:20] += 90.0
    nm, a = orc.search_by_projection_frame(g, desc, None, SF, last, 7.0, 0.0, False, False, True)
    assert nm == n - 20 and np.all(a[:20] == -1) and np.array_equal(a[20:], np.arange(20, n))
    nm, a = orc.search_by_projection_frame(g, desc, None, SF, last, 7.0, 0.0, False, False, False)
    assert nm == n
    # invalid rows, negative depth and out-of-bounds projections are skipped (:2503-2523)
    last["valid"][0] = 0; last["invz"][1] = -1.0; last["u"] = last["u"].copy(); last["u"][2] = 800.0
    nm, a = orc.search_by_projection_frame(g, desc, None, SF, last, 7.0, 0.0, False, False, False)
    assert nm == n - 3 and np.all(a[:3] == -1)


def test_oracle_triangulation_identity(orc, oracle):
    rng = np.random.default_rng(3)
    n = 200
    kps, desc = _unique_frame(orc, rng, n)
    node = (np.arange(n) % 17).astype(np.int32)
    nodes = np.unique(node); start = [0]; idx = []
    for nd in nodes:
        idx += np.nonzero(node == nd)[0].tolist(); start.append(len(idx))
    kf = dict(kps=kps, desc=desc, has_mp=np.zeros(n, np.uint8), uRight=None, fv_node=nodes,
              fv_start=np.array(start, np.int32), fv_idx=np.array(idx, np.int32))
    F = np.zeros((3, 3), np.float32)
    ep = np.array([-1e4, -1e4], np.float32)
    nm, m = orc.search_for_triangulation(oracle, kf, kf, SF, SF * SF, ep, F, False, True, True, 0)   # bCoarse: no epipolar gate
    assert nm == n and np.array_equal(m, np.arange(n))
    nm, m = orc.search_for_triangulation(oracle, kf, kf, SF, SF * SF, ep, F, False, False, True, 0)  # F = 0 -> den == 0 -> rejected
    assert nm == 0 and np.all(m == -1)
    kf2 = dict(kf); kf2["has_mp"] = np.ones(n, np.uint8)
    assert orc.search_for_triangulation(oracle, kf, kf2, SF, SF * SF, ep, F, False, True, True, 0)[0] == 0
    assert orc.search_for_triangulation(oracle, kf, kf, SF, SF * SF, ep, F, True, True, True, 0)[0] == 0   # bOnlyStereo, mono frames


def test_oracle_bow_identity(orc, oracle):
    rng = np.random.default_rng(4)
    n = 200
    kps, desc = _unique_frame(orc, rng, n)
    node = (np.arange(n) % 13).astype(np.int32)
    nodes = np.unique(node); start = [0]; idx = []
    for nd in nodes:
        idx += np.nonzero(node == nd)[0].tolist(); start.append(len(idx))
    kf = dict(kps=kps, desc=desc, valid=np.ones(n, np.uint8), fv_node=nodes, fv_start=np.array(start, np.int32),
              fv_idx=np.array(idx, np.int32))
    for mode in (0, 1):
        nm, m = orc.search_by_bow(oracle, mode, kf, kf, 0.7, True)
        assert nm == n and np.array_equal(m, np.arange(n))
    half = dict(kf); half["valid"] = (np.arange(n) % 2).astype(np.uint8)
    nm, m = orc.search_by_bow(oracle, 0, half, kf, 0.7, True)         # only valid side-1 keypoints are matched
    assert nm == n // 2 and np.array_equal(np.nonzero(m >= 0)[0], np.arange(1, n, 2))
    nm, m = orc.search_by_bow(oracle, 1, kf, half, 0.7, True)         # mode 1 also needs a valid side-2 keypoint
    assert nm == n // 2 and np.all(m[::2] == -1)


def test_oracle_fuse_and_sim3_identity(orc, oracle):
    """Fuse / SearchBySim3 / SearchByProjection(KF, Scw, ...) restatements on a case with a known answer: every
    map point carries the descriptor of exactly one keypoint and projects onto it -> it must pick that keypoint
    (distance 0), independent of the grid walk; a brute-force numpy search gives the same best index."""
    rng = np.random.default_rng(5)
    kps, desc = _unique_frame(orc, rng, 800)
    n = len(kps)
    g = orc.Grid(oracle, kps, 0.0, 752.0, 0.0, 480.0)
    mp = dict(desc=desc.copy(), u=kps["x"].copy(), v=kps["y"].copy(), level=kps["octave"].astype(np.int32), valid=np.ones(n, np.uint8))
    isig2 = (1.0 / (SF * SF)).astype(np.float32)
    inside = (kps["x"] < 752.0) & (kps["y"] < 480.0)
    nf, bi, bd = orc.fuse(g, desc, None, SF, isig2, mp, 3.0)
    assert nf == int(inside.sum()) and np.array_equal(bi[inside], np.arange(n)[inside]) and np.all(bd[inside] == 0)
    mp_ur = dict(mp, ur=(kps["x"] - 5).astype(np.float32))
    uR = (kps["x"] - 5).astype(np.float32)
    nf2, bi2, _ = orc.fuse(g, desc, uR, SF, isig2, mp_ur, 3.0)
    assert nf2 == nf and np.array_equal(bi2, bi)
    # a right coordinate 10 px off fails the 3-dof gate at level 0 (100 / 1 > 7.8) but not at level 7 (100 / 12.8 = 7.8125 > 7.8 too)
    uR_bad = (uR + 10).astype(np.float32)
    nf3, bi3, _ = orc.fuse(g, desc, uR_bad, SF, isig2, mp_ur, 3.0)
    assert nf3 == 0 and np.all(bi3 == -1)
    nm, assign = orc.search_by_projection_sim3(g, desc, SF, mp, 3.0, 1.0, None)
    assert nm == int(inside.sum()) and np.array_equal(assign[inside], np.arange(n)[inside])
    held = np.zeros(n, np.uint8); held[::2] = 1
    nm2, assign2 = orc.search_by_projection_sim3(g, desc, SF, mp, 3.0, 0.0, held)      # only distance 0 passes; held keypoints stay free
    assert np.all(assign2[::2] == -1) and nm2 == int(inside[1::2].sum())
    nf4, m12 = orc.search_by_sim3(g, desc, g, desc, SF, mp, mp, 7.5)
    assert nf4 == int(inside.sum()) and np.array_equal(m12[inside], np.arange(n)[inside])


# ---- the two-camera rig / callback restatements: answers known by construction or from the Nleft == -1 restatement ----

def _one_node_fv(n):
    return dict(fv_node=np.array([3], np.int32), fv_start=np.array([0, n], np.int32), fv_idx=np.arange(n, dtype=np.int32))


def test_oracle_rig_frame_identity_and_empty_left_window(orc, oracle):
    rng = np.random.default_rng(5)
    kL, dL = _unique_frame(orc, rng, 300)
    kR, dR = _unique_frame(orc, rng, 200)
    desc = np.concatenate([dL, dR])
    b = (0.0, 752.0, 0.0, 480.0)
    gL, gR = orc.Grid(oracle, kL, *b), orc.Grid(oracle, kR, *b)
    # rows = the left keypoints themselves, projected exactly; right projections hit right keypoint i % 200
    last = dict(desc=dL, u=kL["x"], v=kL["y"], ur=kR["x"][np.arange(300) % 200], vr=kR["y"][np.arange(300) % 200],
                invz=np.full(300, 0.1, np.float32), octave=np.zeros(300, np.int32), angle=kL["angle"],
                valid=np.ones(300, np.uint8), obs=np.ones(300, np.uint8))
    nm, a = orc.search_by_projection_frame_rig(gL, gR, desc, SF, last, 3.0, False, False, False, None)
    assert np.array_equal(a[:300], np.arange(300))                  # every left keypoint takes its own row
    # random right descriptors: a right keypoint is only accepted when its distance <= TH_HIGH; count consistent
    assert nm == 300 + int(np.sum(a[300:] >= 0))
    # the Nleft == -1 restatement agrees on the left half when there is no right camera
    e = np.zeros(0, orc.KP_DTYPE)
    nm1, a1 = orc.search_by_projection_frame_rig(gL, orc.Grid(oracle, e, *b), dL, SF, last, 3.0, False, False, True, None)
    lastm = {k: v for k, v in last.items() if k not in ("ur", "vr")}
    nm0, a0 = orc.search_by_projection_frame(gL, dL, None, SF, lastm, 3.0, 0.0, False, False, True, None)
    assert nm1 == nm0 and np.array_equal(a1, a0)
    # rows whose LEFT window is empty never reach the right camera (:2541), even with a perfect right match
    last2 = dict(last); last2["desc"] = dR[np.arange(300) % 200]
    last2["u"] = np.full(300, 751.0, np.float32); last2["v"] = np.full(300, 479.0, np.float32)
    nm2, a2 = orc.search_by_projection_frame_rig(gL, gR, desc, SF, last2, 3.0, False, False, False, None)
    assert nm2 == 0 and np.all(a2 == -1)


def test_oracle_rig_bow_quirks(orc, oracle):
    rng = np.random.default_rng(6)
    n = 60
    k1, d1 = _unique_frame(orc, rng, n)
    # Frame: left rows = exact copies of the keyframe descriptors, right rows = copies with 3 flipped bits
    dR = d1.copy(); dR[:, 0] ^= 0x07
    k2 = np.concatenate([k1, k1]); d2 = np.concatenate([d1, dR])
    kf1 = dict(kps=k1, desc=d1, valid=np.ones(n, np.uint8), **_one_node_fv(n))
    kf2 = dict(kps=k2, desc=d2, valid=None, **_one_node_fv(2 * n))
    nm, m = orc.search_by_bow_rig(oracle, kf1, kf2, n, 0.7, False)
    assert nm == 2 * n and np.array_equal(m[:n], np.arange(n)) and np.array_equal(m[n:], np.arange(n))
    # with nleft2 == n2 (no right rows) it is the Nleft == -1 branch
    nm_a, m_a = orc.search_by_bow_rig(oracle, kf1, kf2, 2 * n, 0.7, True)
    nm_b, m_b = orc.search_by_bow(oracle, 0, kf1, kf2, 0.7, True)
    assert nm_a == nm_b and np.array_equal(m_a, m_b)
    # left best fails TH_LOW (random left rows) -> the right match is not even considered (:431 encloses :455)
    d2b = np.concatenate([rng.integers(0, 256, (n, 32), dtype=np.uint8), dR])
    nm_c, m_c = orc.search_by_bow_rig(oracle, kf1, dict(kf2, desc=d2b), n, 0.7, False)
    assert nm_c == 0 and np.all(m_c == -1)
    # right match ignores its own ratio test (`|| true`): two identical right rows per keyframe keypoint still match
    d2c = np.concatenate([d1, dR, dR]); k2c = np.concatenate([k1, k1, k1])
    nm_d, m_d = orc.search_by_bow_rig(oracle, kf1, dict(kps=k2c, desc=d2c, valid=None, **_one_node_fv(3 * n)), n, 0.7, False)
    assert np.array_equal(m_d[n:2 * n], np.arange(n)) and nm_d == 2 * n


def test_oracle_triangulation_callback_order_and_overloads(orc, oracle):
    rng = np.random.default_rng(7)
    n = 80
    k1, d1 = _unique_frame(orc, rng, n)
    d2 = d1.copy(); d2[:, 1] ^= 0x01                               # every pair (i, i) at distance 1, others ~128
    kf1 = dict(kps=k1, desc=d1, has_mp=np.zeros(n, np.uint8), **_one_node_fv(n))
    kf2 = dict(kps=k1, desc=d2, has_mp=np.zeros(n, np.uint8), **_one_node_fv(n))
    calls = []
    nm, m = orc.search_for_triangulation_cb(oracle, 0, kf1, kf2, lambda a, b: calls.append((a, b)) or True, False, False, False)
    assert nm == n and np.array_equal(m, np.arange(n))
    assert calls == [(i, i) for i in range(n)]                     # only pairs with dist <= TH_LOW reach the camera model
    # overload 0 marks vbMatched2: with identical KF2 rows the first idx1 takes idx2 = 0 ... overload 1 / 2 do not mark
    d2s = np.repeat(d1[:1], n, axis=0); d1s = np.repeat(d1[:1], n, axis=0)
    a = dict(kf1, desc=d1s); b = dict(kf2, desc=d2s)
    _, m0 = orc.search_for_triangulation_cb(oracle, 0, a, b, lambda i, j: True, False, False, False)
    _, m1 = orc.search_for_triangulation_cb(oracle, 1, a, b, lambda i, j: True, False, False, False)
    _, m2 = orc.search_for_triangulation_cb(oracle, 2, a, b, lambda i, j: True, False, False, False)
    assert np.array_equal(m0, np.arange(n)[::-1])                 # `dist > bestDist` keeps the LAST equal candidate, then marks it
    assert np.all(m1 == n - 1) and np.all(m2 == n - 1)
    # bOnlyStereo: nothing with a rig (overloads 0, 1); ignored by overload 2
    assert orc.search_for_triangulation_cb(oracle, 0, kf1, kf2, lambda i, j: True, True, False, False)[0] == 0
    assert orc.search_for_triangulation_cb(oracle, 2, kf1, kf2, lambda i, j: True, True, False, False)[0] == n


def test_oracle_fisheye_ratio(orc, oracle):
    rng = np.random.default_rng(8)
    dL = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    dR = np.concatenate([rng.integers(0, 256, (7, 32), dtype=np.uint8), dL[10:]])     # right row 7 + k == left row 10 + k
    good, l2r, r2l, bd = orc.stereo_fisheye_match(oracle, dL, 10, dR, 7)
    assert good == 40 and np.all(l2r[:10] == -1) and np.array_equal(l2r[10:], np.arange(7, 47)) and np.all(bd[10:] == 0)
    assert np.all(r2l[:7] == -1) and np.array_equal(r2l[7:], np.arange(10, 50))
    # fewer than two train rows: knnMatch returns one neighbour, the size() >= 2 test fails
    good, l2r, _, _ = orc.stereo_fisheye_match(oracle, dL, 10, dR[:8], 7)
    assert good == 0 and np.all(l2r == -1)
    # double 0.7: d0 = 7, d1 = 10 -> 7 < 7.000000000000001 is false only if the product rounds to 7.0; pin the arithmetic
    q = np.zeros((1, 32), np.uint8)
    t = np.zeros((2, 32), np.uint8); t[0, 0] = 0x7F; t[1, 0] = 0xFF; t[1, 1] = 0x03
    good, l2r, _, _ = orc.stereo_fisheye_match(oracle, q, 0, t, 0)
    assert good == int(float(7) < float(10) * 0.7)

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

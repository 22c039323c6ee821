"""Second source for the matcher oracles: oracle/_ref/libref_match.so is the UNMODIFIED /root/reference/src/ORBmatcher.cc
compiled against data-holder doubles of Frame / KeyFrame / MapPoint / GeometricCamera (oracle/refshim/ref_stubs.h) and an
OpenCV stand-in (oracle/cvshim).  Every ref_* entry point takes the arguments of the orc_* restatement it pins
(oracle/ref_match_wrap.cc), so each test runs the SAME orc_py wrapper twice - once against liborc.so, once against a
proxy that routes orc_<name> to ref_<name> - and requires identical results.  The GPU is then compared with an oracle that
the reference's own code has confirmed (tests/test_gpu_search.py, test_gpu_rig.py, test_gpu_match.py use the same generators).
"""
import copy
import ctypes as C
import pathlib

import numpy as np
import pytest

from test_gpu_search import SF, SIGMA2, _feature_vector, _flip, _frame, _moved

ROOT = pathlib.Path(__file__).resolve().parents[1]


class _Proxy:
    """looks like the ctypes handle of liborc.so, but orc_<name> resolves to ref_<name> of libref_match.so where that exists"""

    def __init__(self, real, ref):
        self._real, self._ref, self._cache, self.routed = real, ref, {}, set()

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        if name in self._cache:
            return self._cache[name]
        f = getattr(self._real, name)
        if name.startswith("orc_"):
            try:
                g = getattr(self._ref, "ref_" + name[4:])
            except AttributeError:
                g = None
            if g is not None:
                if f.argtypes is not None:
                    g.argtypes = f.argtypes
                g.restype = f.restype
                self.routed.add(name)
                f = g
        self._cache[name] = f
        return f


@pytest.fixture(scope="module")
def refo(orc, oracle):
    path = ROOT / "oracle" / "_ref" / "libref_match.so"
    if not path.exists():
        import ref_py
        try:
            ref_py.build()
        except Exception:
            pass
    if not path.exists():
        pytest.skip("oracle/_ref/libref_match.so not built and /root/reference absent")
    o = copy.copy(oracle)
    o.L = _Proxy(oracle.L, C.CDLL(str(path)))
    return o


def test_descriptor_distance_and_three_maxima(orc, oracle, refo):
    rng = np.random.default_rng(0)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        if rng.random() < 0.2:
            b = a.copy(); b[rng.integers(0, 32)] ^= 1 << rng.integers(0, 8)
        assert oracle.hamming(a, b) == refo.hamming(a, b) == int(np.unpackbits(a ^ b).sum())
    for _ in range(300):
        h = rng.integers(0, rng.choice([2, 5, 50, 500]), 30).astype(np.int32)
        if rng.random() < 0.3:
            h[rng.integers(0, 30)] = 1000
        assert orc.three_maxima(oracle, h) == orc.three_maxima(refo, h)
    assert {"orc_hamming256", "orc_three_maxima"} <= refo.L.routed


def _mp_rows(orc, rng, kps, desc, n, rows, cols):
    return dict(desc=_flip(rng, desc.copy(), 0, 40),
                projx=(kps["x"] + rng.normal(0, 3, n)).astype(np.float32), projy=(kps["y"] + rng.normal(0, 3, n)).astype(np.float32),
                projxr=(kps["x"] - rng.uniform(0, 40, n)).astype(np.float32),
                level=np.clip(kps["octave"] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32),
                viewcos=rng.choice([0.9, 0.999], n).astype(np.float32), depth=rng.uniform(1, 80, n).astype(np.float32),
                inview=(rng.random(n) < 0.9).astype(np.uint8))


@pytest.mark.parametrize("th,stereo,far,seed", [(3.0, False, False, 1), (1.0, True, False, 2), (5.0, True, True, 3), (15.0, False, True, 4)])
def test_search_by_projection_map_points(orc, oracle, refo, th, stereo, far, seed):
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 900
    kps, desc, _, _ = _frame(orc, rng, n, rows, cols)
    mp = _mp_rows(orc, rng, kps, desc, n, rows, cols)
    uR = np.where(rng.random(n) < 0.7, kps["x"] - rng.uniform(1, 40, n), -1).astype(np.float32) if stereo else None
    assign0 = np.where(rng.random(n) < 0.05, rng.integers(0, n, n), -1).astype(np.int32)
    b = (0.0, float(cols), 0.0, float(rows))
    out = []
    for o in (oracle, refo):
        g = orc.Grid(o, kps, *b)
        out.append(orc.search_by_projection(g, desc, uR, SF, mp, th, 0.8, far, 40.0, assign0.copy()))
    assert out[0][0] == out[1][0] > 100 and np.array_equal(out[0][1], out[1][1])
    assert "orc_search_by_projection" in refo.L.routed


@pytest.mark.parametrize("th,far,seed", [(3.0, False, 5), (8.0, True, 6)])
def test_search_by_projection_map_points_rig(orc, oracle, refo, th, far, seed):
    rng = np.random.default_rng(seed)
    rows, cols, nL, nR = 512, 512, 700, 600
    kL, dL, _, _ = _frame(orc, rng, nL, rows, cols)
    kR, dR, _ = _moved(orc, rng, kL[:nR], dL[:nR], rows, cols, shift=(-15.0, 0.5))
    desc = np.concatenate([dL, dR])
    l2r = np.where(rng.random(nL) < 0.3, rng.integers(0, nR, nL), -1).astype(np.int32)
    r2l = np.where(rng.random(nR) < 0.3, rng.integers(0, nL, nR), -1).astype(np.int32)
    n = nL
    mp = dict(desc=_flip(rng, dL.copy(), 0, 40), depth=rng.uniform(1, 80, n).astype(np.float32),
              inviewL=(rng.random(n) < 0.8).astype(np.uint8), projxL=(kL["x"] + rng.normal(0, 3, n)).astype(np.float32),
              projyL=(kL["y"] + rng.normal(0, 3, n)).astype(np.float32), levelL=np.clip(kL["octave"] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32),
              viewcosL=rng.choice([0.9, 0.999], n).astype(np.float32),
              inviewR=(rng.random(n) < 0.8).astype(np.uint8), projxR=(kL["x"] - 15 + rng.normal(0, 3, n)).astype(np.float32),
              projyR=(kL["y"] + 0.5 + rng.normal(0, 3, n)).astype(np.float32),
              levelR=np.where(rng.random(n) < 0.1, -1, np.clip(kL["octave"] + rng.choice([0, 0, 1], n), 0, 7)).astype(np.int32),
              viewcosR=rng.choice([0.9, 0.999], n).astype(np.float32))
    assign0 = np.where(rng.random(nL + nR) < 0.05, rng.integers(0, n, nL + nR), -1).astype(np.int32)
    b = (0.0, float(cols), 0.0, float(rows))
    out = []
    for o in (oracle, refo):
        gL, gR = orc.Grid(o, kL, *b), orc.Grid(o, kR, *b)
        out.append(orc.search_by_projection_rig(gL, gR, desc, l2r, r2l, SF, mp, th, 0.8, far, 40.0, assign0.copy()))
    assert out[0][0] == out[1][0] > 100 and np.array_equal(out[0][1], out[1][1])
    assert "orc_search_by_projection_rig" in refo.L.routed


def _last_rows(orc, rng, lk, ld, n, rig_offset=None):
    # inverse depths that survive z = 1 / invz -> 1 / z exactly (powers of two), a few points behind the camera
    invz = np.where(rng.random(n) < 0.97, 2.0 ** -rng.integers(1, 6, n), -0.1).astype(np.float32)
    last = dict(desc=_flip(rng, ld.copy(), 0, 10),
                u=(lk["x"] + 4.0 + rng.normal(0, 3, n)).astype(np.float32), v=(lk["y"] - 3.0 + rng.normal(0, 3, n)).astype(np.float32),
                invz=invz, octave=lk["octave"].astype(np.int32), angle=lk["angle"],
                valid=(rng.random(n) < 0.85).astype(np.uint8), obs=(rng.random(n) < 0.9).astype(np.uint8))
    if rig_offset is not None:
        last["ur"] = (last["u"] + np.float32(rig_offset[0])).astype(np.float32)
        last["vr"] = (last["v"] + np.float32(rig_offset[1])).astype(np.float32)
    return last


@pytest.mark.parametrize("direction,stereo,orient,seed", [("none", False, True, 11), ("forward", True, True, 12), ("backward", True, False, 13),
                                                         ("none", True, True, 14)])
def test_search_by_projection_frame(orc, oracle, refo, direction, stereo, orient, seed):
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1000
    lk, ld, _, _ = _frame(orc, rng, n, rows, cols)
    ck, cd, _ = _moved(orc, rng, lk, ld, rows, cols)
    uR = np.where(rng.random(n) < 0.7, ck["x"] - rng.uniform(1, 40, n), -1).astype(np.float32) if stereo else None
    last = _last_rows(orc, rng, lk, ld, n)
    held = (rng.random(n) < 0.03).astype(np.uint8)
    th, mbf = (7.0, 40.0) if stereo else (15.0, 0.0)
    fwd, bwd = direction == "forward", direction == "backward"
    b = (0.0, float(cols), 0.0, float(rows))
    out = []
    for o in (oracle, refo):
        g = orc.Grid(o, ck, *b)
        out.append(orc.search_by_projection_frame(g, cd, uR, SF, last, th, mbf, fwd, bwd, orient, held))
    assert out[0][0] == out[1][0] > 100 and np.array_equal(out[0][1], out[1][1])
    assert "orc_search_by_projection_frame" in refo.L.routed


@pytest.mark.parametrize("direction,orient,seed", [("none", True, 21), ("forward", True, 22), ("backward", False, 23)])
def test_search_by_projection_frame_rig(orc, oracle, refo, direction, orient, seed):
    rng = np.random.default_rng(seed)
    rows, cols, nL, nR = 512, 512, 900, 800
    lk, ld, _, _ = _frame(orc, rng, nL, rows, cols)
    ckL, cdL, _ = _moved(orc, rng, lk, ld, rows, cols)
    ckR, cdR, _ = _moved(orc, rng, lk[:nR], ld[:nR], rows, cols, shift=(-16.0, 5.0))
    desc = np.concatenate([cdL, cdR])
    last = _last_rows(orc, rng, lk, ld, nL, rig_offset=(-20.0, 8.0))
    far = rng.random(nL) < 0.08                         # empty LEFT windows: the right half must be skipped too (:2541)
    last["u"][far] = 2.0; last["v"][far] = rows - 1.0
    last["ur"] = (last["u"] + np.float32(-20.0)).astype(np.float32); last["vr"] = (last["v"] + np.float32(8.0)).astype(np.float32)
    held = (rng.random(nL + nR) < 0.03).astype(np.uint8)
    fwd, bwd = direction == "forward", direction == "backward"
    b = (0.0, float(cols), 0.0, float(rows))
    out = []
    for o in (oracle, refo):
        gL, gR = orc.Grid(o, ckL, *b), orc.Grid(o, ckR, *b)
        out.append(orc.search_by_projection_frame_rig(gL, gR, desc, SF, last, 15.0, fwd, bwd, orient, held))
    assert out[1][0] != -1000, "rows do not satisfy the wrapper's constant right-camera offset"
    assert out[0][0] == out[1][0] > 100 and np.array_equal(out[0][1], out[1][1])
    assert int(np.sum(out[0][1][nL:] >= 0)) > 30
    assert "orc_search_by_projection_frame_rig" in refo.L.routed


def _bow_pair(orc, rng, n, flips=(0, 50), nodes=89):
    rows, cols = 480, 752
    k1, d1, pool, src = _frame(orc, rng, n, rows, cols)
    k2, d2, perm = _moved(orc, rng, k1, d1, rows, cols, flips=flips)
    node1 = (src % nodes).astype(np.int32); node2 = node1[perm]
    node2 = np.where(rng.random(n) < 0.05, rng.integers(0, nodes, n), node2).astype(np.int32)
    kf1 = dict(kps=k1, desc=d1); kf2 = dict(kps=k2, desc=d2)
    kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = _feature_vector(rng, node1, 0.02)
    kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = _feature_vector(rng, node2, 0.02)
    return kf1, kf2


@pytest.mark.parametrize("mode,nnratio,orient,seed", [(0, 0.7, True, 31), (0, 0.9, False, 32), (1, 0.8, True, 33), (1, 0.6, True, 34)])
def test_search_by_bow(orc, oracle, refo, mode, nnratio, orient, seed):
    rng = np.random.default_rng(seed)
    n = 1200
    kf1, kf2 = _bow_pair(orc, rng, n)
    kf1["valid"] = (rng.random(n) < 0.7).astype(np.uint8)
    kf2["valid"] = None if mode == 0 else (rng.random(n) < 0.7).astype(np.uint8)
    a = orc.search_by_bow(oracle, mode, kf1, kf2, nnratio, orient)
    b = orc.search_by_bow(refo, mode, kf1, kf2, nnratio, orient)
    if mode == 1:
        # the reference returns KF2's MapPoint* per KF1 keypoint (:1071); the double's `index` is that keypoint's row in KF2
        pass
    assert a[0] == b[0] > 40 and np.array_equal(a[1], b[1])
    assert "orc_search_by_bow" in refo.L.routed


@pytest.mark.parametrize("nnratio,orient,seed,frac", [(0.7, True, 41, 0.55), (0.9, False, 42, 0.5), (0.75, True, 43, 0.0), (1.3, True, 44, 0.4)])
def test_search_by_bow_rig(orc, oracle, refo, nnratio, orient, seed, frac):
    rng = np.random.default_rng(seed)
    n = 1200
    kf1, kf2 = _bow_pair(orc, rng, n)
    kf1["valid"] = (rng.random(n) < 0.7).astype(np.uint8)
    nleft2 = int(n * frac)
    a = orc.search_by_bow_rig(oracle, kf1, kf2, nleft2, nnratio, orient)
    b = orc.search_by_bow_rig(refo, kf1, kf2, nleft2, nnratio, orient)
    assert a[0] == b[0] and np.array_equal(a[1], b[1])
    if nleft2:
        assert a[0] > 40
    assert "orc_search_by_bow_rig" in refo.L.routed


@pytest.mark.parametrize("window,orient,seed", [(100, True, 51), (100, False, 52), (30, True, 53)])
def test_search_for_initialization(orc, oracle, refo, window, orient, seed):
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1500
    k1, d1, _, _ = _frame(orc, rng, n, rows, cols)
    k2, d2, _ = _moved(orc, rng, k1, d1, rows, cols, shift=(12.0, 5.0), flips=(0, 30))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    b = (0.0, float(cols), 0.0, float(rows))
    res = []
    for o in (oracle, refo):
        g2 = orc.Grid(o, k2, *b)
        r1 = orc.search_for_initialization(o, k1, d1, g2, d2, prev, window, 0.9, orient)
        r2 = orc.search_for_initialization(o, k1, d1, g2, d2, r1[2], window, 0.9, orient)     # second round, vbPrevMatched fed back
        res.append((r1, r2))
    for k in (0, 1):
        assert res[0][k][0] == res[1][k][0] and np.array_equal(res[0][k][1], res[1][k][1]) and np.array_equal(res[0][k][2], res[1][k][2])
    assert res[0][0][0] > 50
    assert "orc_search_for_initialization" in refo.L.routed


@pytest.mark.parametrize("variant,only_stereo,coarse,stereo,seed", [(0, False, False, False, 61), (0, False, False, True, 62), (1, False, False, True, 63),
                                                                  (0, True, False, True, 64), (0, False, True, False, 65), (1, False, True, False, 66)])
def test_search_for_triangulation_pinhole(orc, oracle, refo, variant, only_stereo, coarse, stereo, seed):
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1300
    k1, d1, pool, src = _frame(orc, rng, n, rows, cols)
    k2, d2, perm = _moved(orc, rng, k1, d1, rows, cols, shift=(-18.0, 0.0), jitter=0.8, flips=(0, 45))
    fx = fy = 458.0; cx, cy = 367.0, 248.0
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    t12 = np.array([0.11, 0.002, -0.004]); tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    F12 = (np.linalg.inv(K).T @ tx @ np.eye(3) @ np.linalg.inv(K)).astype(np.float32)
    ep = np.array([cx + 120.0, cy - 15.0], np.float32)
    node1 = (src % 97).astype(np.int32); node2 = node1[perm]
    node2 = np.where(rng.random(n) < 0.05, rng.integers(0, 97, n), node2).astype(np.int32)
    kf1 = dict(kps=k1, desc=d1, has_mp=(rng.random(n) < 0.35).astype(np.uint8), uRight=None)
    kf2 = dict(kps=k2, desc=d2, has_mp=(rng.random(n) < 0.35).astype(np.uint8), uRight=None)
    if stereo:
        kf1["uRight"] = np.where(rng.random(n) < 0.5, k1["x"] - 10, -1).astype(np.float32)
        kf2["uRight"] = np.where(rng.random(n) < 0.5, k2["x"] - 10, -1).astype(np.float32)
    kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = _feature_vector(rng, node1, 0.02)
    kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = _feature_vector(rng, node2, 0.02)
    for orient in (True, False):
        a = orc.search_for_triangulation(oracle, kf1, kf2, SF, SIGMA2, ep, F12, only_stereo, coarse, orient, variant)
        b = orc.search_for_triangulation(refo, kf1, kf2, SF, SIGMA2, ep, F12, only_stereo, coarse, orient, variant)
        assert a[0] == b[0] and np.array_equal(a[1], b[1]), f"orient {orient}: {np.sum(a[1] != b[1])} differ"
    assert a[0] > (5 if only_stereo else 30)
    assert "orc_search_for_triangulation" in refo.L.routed


def _verdict_fn(seed, p):
    calls = []

    def v(i1, i2):
        calls.append((i1, i2))
        h = (i1 * 2654435761 + i2 * 40503 + seed * 97) & 0xFFFFFFFF
        h ^= h >> 13; h = (h * 1274126177) & 0xFFFFFFFF; h ^= h >> 16
        return (h % 1000) < int(p * 1000)
    return v, calls


@pytest.mark.parametrize("overload,only_stereo,coarse,orient,p,seed", [(0, False, False, True, 0.6, 71), (1, False, False, True, 0.6, 72),
                                                                      (2, False, False, True, 0.5, 73), (0, False, True, True, 0.5, 74),
                                                                      (1, True, False, True, 0.5, 75), (2, True, True, False, 0.3, 76),
                                                                      (0, False, False, False, 1.0, 77)])
def test_search_for_triangulation_callback(orc, oracle, refo, overload, only_stereo, coarse, orient, p, seed):
    rng = np.random.default_rng(seed)
    n = 1100
    kf1, kf2 = _bow_pair(orc, rng, n, flips=(0, 45), nodes=97)
    kf1["has_mp"] = (rng.random(n) < 0.35).astype(np.uint8)
    kf2["has_mp"] = (rng.random(n) < 0.35).astype(np.uint8)
    va, ca = _verdict_fn(seed, p)
    vb, cb = _verdict_fn(seed, p)
    a = orc.search_for_triangulation_cb(oracle, overload, kf1, kf2, va, only_stereo, coarse, orient)
    b = orc.search_for_triangulation_cb(refo, overload, kf1, kf2, vb, only_stereo, coarse, orient)
    assert a[0] == b[0] and np.array_equal(a[1], b[1]), f"{np.sum(a[1] != b[1])} matches differ"
    assert ca == cb, "the reference asks the camera model for other pairs / in another order"
    assert "orc_search_for_triangulation_cb" in refo.L.routed


# ---- the loop-closing / relocalisation / local-mapping searches (second half) ----

from test_gpu_search import _projected_rows


@pytest.mark.parametrize("th,orb_dist,orient,seed", [(10.0, 100, True, 51), (3.0, 64, True, 52), (10.0, 100, False, 53)])
def test_search_by_projection_keyframe(orc, oracle, refo, th, orb_dist, orient, seed):
    """ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:2723-2862 (Relocalization)."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1500
    kkps, kdesc, _, _ = _frame(orc, rng, n, rows, cols)
    ckps, cdesc, perm = _moved(orc, rng, kkps, kdesc, rows, cols)
    mp = dict(desc=_flip(rng, kdesc.copy(), 0, 10),
              u=(kkps["x"] + 4.0 + rng.normal(0, 2, n)).astype(np.float32), v=(kkps["y"] - 3.0 + rng.normal(0, 2, n)).astype(np.float32),
              level=np.clip(kkps["octave"] + rng.choice([-1, 0, 0, 1], n), 0, 7).astype(np.int32), angle=kkps["angle"],
              valid=(rng.random(n) < 0.8).astype(np.uint8))
    mp["u"][:4] = [0.0, float(cols), -0.5, float(cols) + 0.5]        # the frame's bounds are inclusive here (:2761-2764)
    held = (rng.random(n) < 0.1).astype(np.uint8)
    b = (0.0, float(cols), 0.0, float(rows))
    out = [orc.search_by_projection_keyframe(orc.Grid(o, ckps, *b), cdesc, SF, mp, th, orb_dist, orient, held) for o in (oracle, refo)]
    assert out[0][0] == out[1][0] > 100 and np.array_equal(out[0][1], out[1][1])
    assert "orc_search_by_projection_keyframe" in refo.L.routed


@pytest.mark.parametrize("th,ratio,seed,use_held", [(3.0, 1.0, 61, True), (8.0, 1.5, 62, True), (5.0, 0.6, 63, False), (8.0, 5.0, 64, True),
                                                    (3.0, 1.0, 65, True)])
def test_search_by_projection_sim3(orc, oracle, refo, th, ratio, seed, use_held):
    """ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:549-678, and the
    vpPointsKFs / vpMatchedKF twin :681-797 (the wrapper alternates between the two overloads from call to call).
    ratioHamming stays below 256 / TH_LOW = 5.12: above it a map point without any candidate at its level passes
    `bestDist <= TH_LOW * ratioHamming` with bestIdx = -1 and the reference writes vpMatched[-1] (:667-670; it aborts here). The
    restatement and the product skip such a row (tests/test_gpu_search.py runs ratio 6.0)."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1500
    kkps, kdesc, _, _ = _frame(orc, rng, n, rows, cols)
    ckps, cdesc, perm = _moved(orc, rng, kkps, kdesc, rows, cols)
    mp = _projected_rows(orc, rng, kkps, kdesc, n, rows, cols)
    held = (rng.random(n) < 0.15).astype(np.uint8) if use_held else None
    b = (0.0, float(cols), 0.0, float(rows))
    out = [orc.search_by_projection_sim3(orc.Grid(o, ckps, *b), cdesc, SF, mp, th, ratio, held) for o in (oracle, refo)]
    assert out[0][0] == out[1][0] > 50 and np.array_equal(out[0][1], out[1][1])
    assert "orc_search_by_projection_sim3" in refo.L.routed


def _exact_disparity_rows(rng, mp, n):
    """ur = u - d with d a power of two: the reference recomputes ur = uv.x - bf * (1 / z) from the depth (:1912), and the
    double's z = 1 / d must reproduce ur in float arithmetic; rows where it would not are taken out (valid = 0)."""
    d = (2.0 ** rng.integers(-2, 6, n)).astype(np.float32)
    u = mp["u"].astype(np.float32)
    ur = (u - d).astype(np.float32)
    dd = (u - ur).astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        z = (np.float32(1) / dd).astype(np.float32)
        back = (u - (np.float32(1) / z).astype(np.float32)).astype(np.float32)
    ok = (dd > 0) & (back == ur)
    mp["ur"] = ur
    mp["valid"] = (mp["valid"].astype(bool) & ok).astype(np.uint8)
    return mp


@pytest.mark.parametrize("form,stereo,th,seed", [("pose", True, 3.0, 71), ("pose", False, 3.0, 72), ("sim3", False, 4.0, 73), ("pose", True, 10.0, 74),
                                                 ("sim3", False, 10.0, 75)])
def test_fuse(orc, oracle, refo, form, stereo, th, seed):
    """ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th, bRight) src/ORBmatcher.cc:1823-2035 and Fuse(KeyFrame*, Scw, vpPoints, th,
    vpReplacePoint) :2051-2185.  The fuse target of every row is read back from what the reference did to the map
    (AddObservation / Replace / vpReplacePoint); the reference does not expose bestDist, so distances are checked on the oracle's side."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1800
    kkps, kdesc, _, _ = _frame(orc, rng, n, rows, cols, quant=0.5)
    ckps, cdesc, perm = _moved(orc, rng, kkps, kdesc, rows, cols)
    mp = _projected_rows(orc, rng, kkps, kdesc, n, rows, cols)
    if form == "pose":
        mp = _exact_disparity_rows(rng, mp, n)
    src, dst = rng.integers(0, n, 60), rng.integers(0, n, 60)      # twin rows: two map points that fuse with the same keypoint
    for k in mp:
        mp[k][dst] = mp[k][src]
    uR = None
    if stereo:
        uR = np.where(rng.random(n) < 0.6, ckps["x"] - rng.uniform(0, 40, n), -1.0).astype(np.float32)
    b = (0.0, float(cols), 0.0, float(rows))
    isig2 = (1.0 / SIGMA2).astype(np.float32)
    nf_o, bi_o, bd_o = orc.fuse(orc.Grid(oracle, ckps, *b), cdesc, uR, SF, isig2, mp, th)
    nf_r, bi_r, _ = orc.fuse(orc.Grid(refo, ckps, *b), cdesc, uR, SF, isig2, mp, th)
    assert nf_r != -1000
    assert nf_o == nf_r > 30
    assert np.array_equal(bi_o, bi_r), f"{np.sum(bi_o != bi_r)} fuse targets differ"
    # several rows land on the same keypoint (the Replace branch of the reference): make sure the case occurred
    hit = bi_o[bi_o >= 0]
    assert len(hit) > len(np.unique(hit))
    for i in np.flatnonzero(bi_o >= 0)[:200]:
        assert bd_o[i] == int(np.unpackbits(mp["desc"][i] ^ cdesc[bi_o[i]]).sum()) <= 50
    assert "orc_fuse" in refo.L.routed


@pytest.mark.parametrize("th,seed", [(7.5, 81), (3.0, 82), (7.5, 83)])
def test_search_by_sim3(orc, oracle, refo, th, seed):
    """ORBmatcher::SearchBySim3, src/ORBmatcher.cc:2201-2434 (both projection directions and the cross-check)."""
    rng = np.random.default_rng(seed)
    rows, cols, n = 480, 752, 1400
    k1, d1, _, _ = _frame(orc, rng, n, rows, cols)
    k2, d2, perm = _moved(orc, rng, k1, d1, rows, cols, flips=(0, 25))
    inv = np.empty(n, np.int64); inv[perm] = np.arange(n)
    rows1 = dict(desc=_flip(rng, d1.copy(), 0, 10), u=(k2["x"][inv] + rng.normal(0, 1.5, n)).astype(np.float32),
                 v=(k2["y"][inv] + rng.normal(0, 1.5, n)).astype(np.float32),
                 level=np.clip(k2["octave"][inv] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32), valid=(rng.random(n) < 0.8).astype(np.uint8))
    rows2 = dict(desc=_flip(rng, d2.copy(), 0, 10), u=(k1["x"][perm] + rng.normal(0, 1.5, n)).astype(np.float32),
                 v=(k1["y"][perm] + rng.normal(0, 1.5, n)).astype(np.float32),
                 level=np.clip(k1["octave"][perm] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32), valid=(rng.random(n) < 0.8).astype(np.uint8))
    rows1["u"][:3] = [0.0, float(cols), -0.5]
    b = (0.0, float(cols), 0.0, float(rows))
    out = [orc.search_by_sim3(orc.Grid(o, k1, *b), d1, orc.Grid(o, k2, *b), d2, SF, rows1, rows2, th) for o in (oracle, refo)]
    assert out[0][0] == out[1][0] > 100 and np.array_equal(out[0][1], out[1][1])
    assert "orc_search_by_sim3" in refo.L.routed

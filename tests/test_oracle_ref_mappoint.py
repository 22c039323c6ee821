"""Second source for the MapPoint.cc rows: oracle/_ref/libref_mp.so is the UNMODIFIED /root/reference/src/MapPoint.cc compiled with
the real include/MapPoint.h (doubles for KeyFrame / Frame / Map, oracle/refshim_mp/).
  MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:450-531) == orc_distinctive_descriptors
  MapPoint::PredictScale (:698-741)                                == the level formula of orc_is_in_frustum
"""
import ctypes as C
import pathlib

import numpy as np
import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
VP = C.c_void_p


@pytest.fixture(scope="module")
def refmp():
    path = ROOT / "oracle" / "_ref" / "libref_mp.so"
    if not path.exists():
        import ref_py
        try:
            ref_py.build()
        except Exception:
            pass
    if not path.exists():
        pytest.skip("oracle/_ref/libref_mp.so not built and /root/reference absent")
    L = C.CDLL(str(path))
    L.ref_mp_distinctive.argtypes = [C.c_int, VP, VP, C.c_int, VP, VP]
    L.ref_mp_predict_scale.argtypes = [C.c_int, VP, VP, C.c_float, C.c_int, VP]
    return L


def _groups(rng, maxn, npts):
    sizes = rng.integers(0, maxn + 1, npts)
    sizes[:4] = [0, 1, 2, maxn]
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = rng.integers(0, 256, (npts, 32), dtype=np.uint8)
    desc = np.repeat(base, sizes, axis=0)
    bits = np.unpackbits(desc, axis=1)
    flips = rng.random(bits.shape) < rng.choice([0.0, 0.02, 0.1], (len(bits), 1))      # ties are frequent
    return start, np.ascontiguousarray(np.packbits(bits ^ flips.astype(np.uint8), axis=1))


@pytest.mark.parametrize("seed,maxn,pairs", [(1, 12, 0), (2, 40, 0), (3, 300, 0), (4, 25, 1), (5, 7, 1)])
def test_distinctive_descriptors_equal_reference(orc, oracle, refmp, seed, maxn, pairs):
    """pairs = 1: consecutive descriptors are the left and right observation of one rig keyframe (NLeft != -1, :479-486) - the
    reference then visits them in the same flat order the CSR group lists them."""
    rng = np.random.default_rng(seed)
    npts = 500 if maxn < 100 else 40
    start, desc = _groups(rng, maxn, npts)
    bi_o, bm_o = orc.distinctive_descriptors(oracle, start, desc)
    out = np.empty((npts, 32), np.uint8); has = np.empty(npts, np.uint8)
    refmp.ref_mp_distinctive(npts, start.ctypes.data, desc.ctypes.data, pairs, out.ctypes.data, has.ctypes.data)
    assert np.array_equal(has != 0, bi_o >= 0)
    ties = 0
    for g in np.flatnonzero(bi_o >= 0):
        rows = desc[start[g]:start[g + 1]]
        assert np.array_equal(out[g], rows[bi_o[g]]), g
        # the chosen row is the FIRST of the rows with the smallest median (strict <, :520): check the restatement's median too
        d = np.unpackbits(rows[:, None, :] ^ rows[None, :, :], axis=2).sum(2)
        med = np.sort(d, axis=1)[:, int(0.5 * (len(rows) - 1))]
        assert bm_o[g] == med.min() and bi_o[g] == int(np.argmin(med))
        ties += int((med == med.min()).sum() > 1)
    assert ties > 20


def test_predict_scale_equals_reference(orc, oracle, refmp):
    """ceil(log(mfMaxDistance / dist) / mfLogScaleFactor) clamped to [0, nLevels - 1], ratios on and one ulp around powers of 1.2"""
    rng = np.random.default_rng(7)
    logsf = float(np.log(np.float32(1.2)))
    sf = np.float32(1.2) ** np.arange(-2, 10, dtype=np.float32)
    ratios = np.concatenate([sf, np.nextafter(sf, np.float32(0)), np.nextafter(sf, np.float32(100)), rng.uniform(0.3, 6.0, 4000).astype(np.float32)])
    dist = rng.uniform(0.5, 40.0, len(ratios)).astype(np.float32)
    maxd = (ratios * dist).astype(np.float32)
    lvl = np.empty(len(dist), np.int32)
    assert refmp.ref_mp_predict_scale(len(dist), maxd.ctypes.data, dist.ctypes.data, logsf, 8, lvl.ctypes.data) == 0
    # the oracle's isInFrustum with a point straight ahead at distance dist: level = PredictScale
    m = len(dist)
    pos = np.zeros((m, 3), np.float32); pos[:, 2] = dist
    pts = dict(pos=pos, normal=np.tile(np.array([0, 0, 1], np.float32), (m, 1)), min_dist=np.zeros(m, np.float32), max_dist=maxd,
               desc=np.zeros((m, 32), np.uint8), candidate=np.ones(m, np.uint8))
    pose = dict(Rcw=np.eye(3, dtype=np.float32), tcw=np.zeros(3, np.float32), Ow=np.zeros(3, np.float32))
    cam = dict(fx=400.0, fy=400.0, cx=320.0, cy=240.0, mbf=40.0)
    _, o = orc.is_in_frustum(oracle, pose, cam, (0.0, 640.0, 0.0, 480.0), logsf, 8, pts, 0.5)
    v = o["inview"] != 0
    assert v.sum() > 3000                                   # dist <= 1.2 * maxd for all but the smallest ratios
    assert np.array_equal(o["level"][v], lvl[v])
    assert set(np.unique(lvl)) == set(range(8))

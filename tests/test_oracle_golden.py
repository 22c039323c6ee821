"""CPU suite: the dependency-free C oracle (oracle/orc_*.c) against the committed golden vectors that
oracle/make_golden.py produced with the REAL OpenCV primitives (cv2 4.13.0).  This is what pins the
oracle; the GPU parity tests then compare the CUDA path with the oracle."""
import hashlib
import pathlib

import numpy as np
import pytest

GOLD = pathlib.Path(__file__).resolve().parent / "golden"


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def cand_arr(c):
    return np.stack([c["x"], c["y"], c["score"]], axis=1).astype(np.int32) if len(c) else np.empty((0, 3), np.int32)


@pytest.mark.parametrize("name", ["c1_mono", "c3_fisheye_partial", "c4_kitti", "small"])
def test_extract_matches_cv2_golden(pkg_synth, orc, oracle, name):
    g = np.load(GOLD / f"extract_{name}.npz")
    rows, cols, nf, seed = int(g["rows"]), int(g["cols"]), int(g["nfeatures"]), int(g["seed"])
    img = pkg_synth.synth_frame(rows, cols, seed)
    assert sha(img) == str(g["input_sha"]), "synthetic generator drifted: regenerate goldens"
    E = orc.Extractor(nf, oracle=oracle)
    mono, kps, desc = E(img, tuple(int(v) for v in g["lap"]))
    # constructor tables (src/ORBextractor.cc:468-571)
    assert np.array_equal(E.scale_factors, g["scale_factors"]) and np.array_equal(E.inv_scale_factors, g["inv_scale_factors"])
    assert np.array_equal(E.level_sigma2, g["sigma2"]) and np.array_equal(E.inv_level_sigma2, g["inv_sigma2"])
    assert np.array_equal(E.features_per_level, g["nper"]) and np.array_equal(E.umax, g["umax"])
    off = 0
    for l in range(8):
        assert sha(E.level_plane(l)) == str(g["plane_sha"][l]), f"pyramid level {l}"
        c = cand_arr(E.level_candidates(l))
        assert len(c) == g["ncand"][l] and sha(c) == str(g["cand_sha"][l]), f"FAST candidates level {l}"
        s, a = E.level_selected(l)
        n = int(g["nsel"][l])
        assert np.array_equal(cand_arr(s), g["sel"][off:off + n]), f"quadtree level {l}"
        assert np.array_equal(a, g["angles"][off:off + n]), f"angles level {l}"
        off += n
        b = E.level_blurred(l)
        if b is not None:
            assert sha(b) == str(g["blur_sha"][l]), f"blur level {l}"
    assert mono == int(g["mono"])
    assert kps.tobytes() == g["kps"].tobytes()
    assert np.array_equal(desc, g["desc"])
    # the goldens exercise the minThFAST retry and empty cells
    assert g["cell_stats"][0][1] > 0 and g["cell_stats"][0][2] > 0


def test_stereo_matches_cv2_golden(pkg_synth, orc, oracle):
    g = np.load(GOLD / "stereo_c2.npz")
    c = pkg_synth.CONFIGS["C2"]
    l, r = pkg_synth.synth_stereo_pair(c["rows"], c["cols"], int(g["seed"]))
    assert sha(l) == str(g["left_sha"]) and sha(r) == str(g["right_sha"])
    EL, ER = orc.Extractor(c["nfeatures"], oracle=oracle), orc.Extractor(c["nfeatures"], oracle=oracle)
    _, kl, dl = EL(l); _, kr, dr = ER(r)
    assert sha(kl) == str(g["kps_left_sha"]) and sha(kr) == str(g["kps_right_sha"])
    uR, depth, _, _ = orc.stereo_match(EL, ER, kl, dl, kr, dr, float(g["mb"]), float(g["mbf"]))
    assert np.array_equal(uR, g["uRight"]) and np.array_equal(depth, g["depth"])
    assert (uR >= 0).sum() > 100


def test_primitives_match_cv2_golden(oracle):
    g = np.load(GOLD / "primitives.npz")
    L = oracle.L
    at = np.array([L.orc_fast_atan2(float(y), float(x)) for y, x in zip(g["atan_y"], g["atan_x"])], np.float32)
    assert np.array_equal(at, g["atan"])
    img = g["img"]
    assert np.array_equal(oracle.resize(img, 100, 133), g["resize_133x100"])
    assert np.array_equal(oracle.gaussian(img), g["blur"])
    assert np.array_equal(oracle.border101(img, 19), g["border"])
    for t in (20, 7):
        assert np.array_equal(cand_arr(oracle.fast(img, t)), g[f"fast{t}"])
    idx, dist = oracle.knn2(g["knn_q"], g["knn_t"])
    assert np.array_equal(idx, g["knn_idx"]) and np.array_equal(dist, g["knn_dist"])


def test_octree_tie_break_sensitivity(pkg_synth, orc, oracle):
    """The reference orders equal-size nodes by heap pointer (src/ORBextractor.cc:927); the canonical rule
    is (size, creation order).  Reversing the tie rule must change only a small fraction of keypoints."""
    img = pkg_synth.synth_frame(480, 752, 0)
    a = orc.Extractor(1000, oracle=oracle, tie_mode=0)(img)[1]
    b = orc.Extractor(1000, oracle=oracle, tie_mode=1)(img)[1]
    sa = set(zip(a["x"].tolist(), a["y"].tolist(), a["octave"].tolist()))
    sb = set(zip(b["x"].tolist(), b["y"].tolist(), b["octave"].tolist()))
    frac = len(sa ^ sb) / 2 / len(sa)
    assert 0 < frac < 0.05, frac


def test_trig_and_fma_canonicalisation(pkg_synth, orc):
    """orc_sinf/orc_cosf (glibc algorithm, no FMA) equal the host libm on a dense sample of [0, 2*pi];
    the -ffp-contract=fast -mfma build of the oracle changes < 0.1% of descriptors (north_star bar)."""
    import ctypes, ctypes.util
    o = orc.Oracle()
    libm = ctypes.CDLL(ctypes.util.find_library("m"))
    libm.sinf.restype = libm.cosf.restype = ctypes.c_float
    libm.sinf.argtypes = libm.cosf.argtypes = [ctypes.c_float]
    xs = np.linspace(0, 6.3, 20001, dtype=np.float32)
    for x in xs:
        assert o.L.orc_sinf(float(x)) == libm.sinf(float(x)) and o.L.orc_cosf(float(x)) == libm.cosf(float(x))
    img = pkg_synth.synth_frame(480, 752, 2)
    d0 = orc.Extractor(1000, oracle=o)(img)[2]
    d1 = orc.Extractor(1000, oracle=orc.Oracle(fma=True))(img)[2]
    assert d0.shape == d1.shape
    assert (d0 != d1).any(axis=1).mean() < 0.001

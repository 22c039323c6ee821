"""CPU suite: the oracle's OpenCV-primitive restatements against the live cv2 wheel (4.13.0 in this
image) on fresh random inputs — resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2, BFMatcher —
plus one full operator() against oracle A (oracle/ref_cv2.py, real cv2 calls)."""
import sys
import pathlib

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
ROOT = pathlib.Path(__file__).resolve().parents[1]


def cand_arr(c):
    return np.stack([c["x"], c["y"], c["score"]], axis=1).astype(np.int32) if len(c) else np.empty((0, 3), np.int32)


@pytest.mark.parametrize("shape,seed", [((97, 131), 1), ((240, 376), 2), ((64, 64), 3)])
def test_pixel_primitives(pkg_synth, oracle, shape, seed):
    rng = np.random.default_rng(seed)
    img = pkg_synth.synth_frame(shape[0], shape[1], seed) if shape[0] > 64 else rng.integers(0, 256, shape, dtype=np.uint8)
    for s in (1.2, 1.44, 2.5):
        dh, dw = int(round(shape[0] / s)), int(round(shape[1] / s))
        assert np.array_equal(oracle.resize(img, dh, dw), cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR))
    assert np.array_equal(oracle.border101(img, 19), cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))
    assert np.array_equal(oracle.gaussian(img), cv2.GaussianBlur(img, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101))
    for t in (20, 7):
        k = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(img, None)
        ref = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in k], np.int32).reshape(-1, 3)
        assert np.array_equal(cand_arr(oracle.fast(img, t)), ref)


def test_fast_on_tiny_and_degenerate_images(oracle):
    for shape in [(6, 40), (40, 6), (7, 7), (8, 9)]:
        img = np.random.default_rng(1).integers(0, 256, shape, dtype=np.uint8)
        k = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(img, None)
        assert len(oracle.fast(img, 7)) == len(k)


def test_fast_atan2(oracle):
    rng = np.random.default_rng(0)
    ys = rng.integers(-1_900_000, 1_900_000, 3000).astype(np.float32)
    xs = rng.integers(-1_900_000, 1_900_000, 3000).astype(np.float32)
    for y, x in zip(ys, xs):
        assert oracle.L.orc_fast_atan2(float(y), float(x)) == np.float32(cv2.fastAtan2(float(y), float(x)))


def test_bf_knn(oracle):
    rng = np.random.default_rng(5)
    q = rng.integers(0, 256, (90, 32), dtype=np.uint8); t = rng.integers(0, 256, (70, 32), dtype=np.uint8)
    t[10] = t[60]; q[5] = t[10]
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    idx, dist = oracle.knn2(q, t)
    assert np.array_equal(idx, np.array([[a.trainIdx, b.trainIdx] for a, b in m], np.int32))
    assert np.array_equal(dist, np.array([[int(a.distance), int(b.distance)] for a, b in m], np.int32))


def test_full_operator_against_cv2_oracle(pkg_synth, orc, oracle):
    sys.path.insert(0, str(ROOT / "oracle"))
    import ref_cv2
    img = pkg_synth.synth_frame(300, 400, 21)
    P = ref_cv2.Params(600)
    R = ref_cv2.extract(img, P, (100, 250))
    mono, kps, desc = orc.Extractor(600, oracle=oracle)(img, (100, 250))
    assert mono == R["mono"] and kps.tobytes() == R["kps"].tobytes() and np.array_equal(desc, R["desc"])


def rectify_maps(rows, cols, seed=0):
    """EuRoC-like undistort+rectify maps through the real cv2.initUndistortRectifyMap (stereo_euroc.cc:117)."""
    K = np.array([[0.61 * cols, 0, 0.49 * cols], [0, 0.95 * rows, 0.52 * rows], [0, 0, 1]], np.float64)
    D = np.array([-0.2834, 0.0739, 0.00019, 1.76e-05, 0.0], np.float64)
    rng = np.random.default_rng(seed)
    rv = rng.normal(0, 0.01, 3)
    R, _ = cv2.Rodrigues(rv)
    P = K.copy(); P[0, 0] *= 0.95; P[1, 1] *= 0.95
    return cv2.initUndistortRectifyMap(K, D, R, P, (cols, rows), cv2.CV_32F)


@pytest.mark.parametrize("channels", [1, 3, 4])
def test_prepare_primitives(oracle, channels):
    """remap(INTER_LINEAR, CV_32FC1 maps) and cvtColor(..2GRAY) restatements, incl. taps that leave the source."""
    rng = np.random.default_rng(10 + channels)
    H, W = 120, 168
    shape = (H, W) if channels == 1 else (H, W, channels)
    src = rng.integers(0, 256, shape, dtype=np.uint8)
    m1, m2 = rectify_maps(H, W)
    assert np.array_equal(oracle.remap(src, m1, m2), cv2.remap(src, m1, m2, cv2.INTER_LINEAR))
    # maps of a different size than the source, wild coordinates, exact half fractions
    mx = rng.uniform(-4, W + 4, (77, 93)).astype(np.float32)
    my = rng.uniform(-4, H + 4, (77, 93)).astype(np.float32)
    mx[0, :6] = [-0.5, -1.0, -1.015625, W - 1, W - 0.5, 0.515625]
    my[0, :6] = [0.0, 0.5, H - 1, H - 0.984375, -0.015625, 0.015625]
    assert np.array_equal(oracle.remap(src, mx, my), cv2.remap(src, mx, my, cv2.INTER_LINEAR))
    if channels > 1:
        codes = {(3, True): cv2.COLOR_RGB2GRAY, (3, False): cv2.COLOR_BGR2GRAY,
                 (4, True): cv2.COLOR_RGBA2GRAY, (4, False): cv2.COLOR_BGRA2GRAY}
        for rgb in (True, False):
            assert np.array_equal(oracle.gray(src, rgb), cv2.cvtColor(src, codes[(channels, rgb)]))
            ref = cv2.cvtColor(cv2.remap(src, m1, m2, cv2.INTER_LINEAR), codes[(channels, rgb)])
            assert np.array_equal(oracle.prepare(src, m1, m2, rgb), ref)


def test_undistort_points(oracle):
    """cv::undistortPoints(src, dst, K, dist, Mat(), K) as Frame::UndistortKeyPoints calls it (src/Frame.cc:938): the
    restatement of cvUndistortPointsInternal equals the wheel bit for bit (4- and 5-coefficient models, incl. points
    outside the image and the image corners of ComputeImageBounds)."""
    rng = np.random.default_rng(0)
    for case in range(4):
        fx, fy = np.float32(rng.uniform(180, 900)), np.float32(rng.uniform(180, 900))
        cx, cy = np.float32(rng.uniform(300, 400)), np.float32(rng.uniform(200, 280))
        dist = np.array([rng.uniform(-0.35, 0.1), rng.uniform(-0.1, 0.15), rng.uniform(-2e-3, 2e-3), rng.uniform(-2e-3, 2e-3),
                         rng.uniform(-0.05, 0.05)], np.float32)
        if case % 2 == 0:
            dist = dist[:4]
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        xy = np.stack([rng.uniform(-20, 772, 50000), rng.uniform(-20, 500, 50000)], 1).astype(np.float32)
        xy[:4] = [[0, 0], [752, 0], [0, 480], [752, 480]]
        ref = cv2.undistortPoints(xy.reshape(-1, 1, 2), K, dist, None, K).reshape(-1, 2)
        got = oracle.undistort_points(xy, float(fx), float(fy), float(cx), float(cy), dist)
        assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))

"""GPU parity of the frame preparation in front of operator(): cv::remap(INTER_LINEAR,
CV_32FC1 maps) + cvtColor(..2GRAY) fused on the device, against the C oracle (orc_prepare.c, pinned to
cv2 4.13 in tests/test_oracle_vs_cv2.py).  Bar: prepared pixels bit-exact; extraction on them bit-exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [
    # name, rect rows, rect cols, src rows, src cols, channels, rgb, maps
    ("euroc-gray-rectify", 480, 752, 480, 752, 1, False, True),
    ("bgr-no-maps", 480, 640, 480, 640, 3, False, False),
    ("rgb-no-maps", 480, 640, 480, 640, 3, True, False),
    ("bgra-no-maps", 376, 1241, 376, 1241, 4, False, False),
    ("rgb-rectify", 480, 752, 480, 752, 3, True, True),
    ("rgba-rectify-resize", 400, 640, 480, 752, 4, True, True),
    ("kitti-odd-width-rectify", 376, 1241, 376, 1241, 1, False, True),
    ("odd-width-bgr-rectify", 300, 501, 310, 517, 3, False, True),
]


def _raw(pkg, rows, cols, ch, seed):
    return pkg.synth.synth_frame(rows, cols, seed) if ch == 1 else pkg.synth.synth_color_frame(rows, cols, seed, ch)


@pytest.mark.parametrize("name,rows,cols,srows,scols,ch,rgb,maps", CASES)
def test_prepare_extract_matches_oracle(pkg, orc, oracle, name, rows, cols, srows, scols, ch, rgb, maps):
    raw = _raw(pkg, srows, scols, ch, 31)
    mx = my = None
    X = pkg.ORBextractor(800)
    if maps:
        mx, my = pkg.synth.synth_rectify_maps(rows, cols, srows, scols, seed=3)
        X.set_rectify_maps(mx, my)
    want = oracle.prepare(raw, mx, my, rgb)
    assert want.shape == (rows, cols)
    mono, kps, desc = X.prepare_extract(raw, (0, 0), rgb=rgb)
    got = X.prepared(0)
    assert np.array_equal(got, want), f"{name}: {(got != want).sum()} prepared pixels differ"
    assert np.array_equal(X.pyramid_level(0), want)
    E = orc.Extractor(800, oracle=oracle)
    mono_o, kps_o, desc_o = E(want, (0, 0))
    assert mono == mono_o and np.array_equal(kps, kps_o) and np.array_equal(desc, desc_o)
    X.close()


def test_maps_leave_the_source(pkg, oracle):
    """maps far outside the source (all four taps read the constant border) and exactly on its edges"""
    rows, cols = 200, 240
    raw = pkg.synth.synth_frame(rows, cols, 2, bands=False)
    rng = np.random.default_rng(9)
    mx = rng.uniform(-40, cols + 40, (rows, cols)).astype(np.float32)
    my = rng.uniform(-40, rows + 40, (rows, cols)).astype(np.float32)
    mx[0, :8] = [-1.0, -0.5, -0.03125, 0.0, cols - 1, cols - 1 + 0.5, cols - 0.015625, cols]
    my[0, :8] = [0.0, rows - 1, rows - 0.5, -1.0, -0.984375, rows, 5.0, 5.0]
    mx[1, :4] = [1e9, -1e9, 3.0e4, -3.3e4]
    X = pkg.ORBextractor(100, nlevels=4)
    X.set_rectify_maps(mx, my)
    X.prepare_upload(raw[None])
    assert np.array_equal(X.prepared(0), oracle.remap(raw, mx, my))
    X.close()


def test_prepare_batch_ring_and_resident(pkg, orc, oracle, monkeypatch):
    """batch API through the chunked H2D / prepare / extract / D2H ring (several chunks), and the resident twins"""
    monkeypatch.setenv("ORBX_CHUNK", "2")
    rows, cols, B = 240, 376, 7
    raws = np.stack([pkg.synth.synth_color_frame(rows, cols, 50 + i, 3) for i in range(B)])
    mx, my = pkg.synth.synth_rectify_maps(rows, cols, seed=1)
    X = pkg.ORBextractor(400)
    X.reserve(rows, cols, 4)
    X.set_rectify_maps(mx, my)
    mono, nkp, kps, desc = X.prepare_extract_batch(raws, (0, 0), rgb=False)
    E = orc.Extractor(400, oracle=oracle)
    for f in range(B):
        want = oracle.prepare(raws[f], mx, my, False)
        m_o, k_o, d_o = E(want, (0, 0))
        assert mono[f] == m_o and nkp[f] == len(k_o), f
        assert np.array_equal(kps[f, :nkp[f]], k_o) and np.array_equal(desc[f, :nkp[f]], d_o), f
    # resident: upload + prepare once, run, download
    X.prepare_upload(raws[:4])
    X.prepare_resident(); X.run_resident((0, 0))
    mono2, nkp2, kps2, desc2 = X.download()
    assert np.array_equal(nkp2, nkp[:4]) and np.array_equal(kps2[0, :nkp2[0]], kps[0, :nkp[0]])
    for f in range(4):
        assert np.array_equal(X.prepared(f), oracle.prepare(raws[f], mx, my, False))
    # clearing the maps: colour conversion only
    X.set_rectify_maps(None, None)
    m3, k3, d3 = X.prepare_extract(raws[0])
    assert np.array_equal(X.prepared(0), oracle.gray(raws[0], False))
    # gray frame without maps: plain operator()
    g = oracle.gray(raws[0], False)
    m4, k4, d4 = X.prepare_extract(g)
    assert m4 == m3 and np.array_equal(k4, k3) and np.array_equal(d4, d3)
    X.close()


def test_prepare_argument_errors(pkg):
    X = pkg.ORBextractor(100)
    lib = pkg.binding.load()
    img = np.zeros((64, 64, 2), np.uint8)
    kp = np.empty(10, pkg.KP_DTYPE); d = np.empty((10, 32), np.uint8)
    import ctypes as C
    n = C.c_int(0)
    assert lib.orbx_prepare_extract(X.h, img.ctypes.data, 64, 64, 2, 0, 128, 0, 0, kp.ctypes.data, d.ctypes.data, 10, C.byref(n)) == pkg.binding.ORBX_ERR_ARG
    assert lib.orbx_prepare_extract(X.h, None, 0, 0, 1, 0, 0, 0, 0, kp.ctypes.data, d.ctypes.data, 10, C.byref(n)) == pkg.binding.ORBX_EMPTY_IMAGE
    mx = np.zeros((8, 8), np.float32)
    assert lib.orbx_set_rectify_maps(X.h, mx.ctypes.data, None, 8, 8, 32) == pkg.binding.ORBX_ERR_ARG
    assert lib.orbx_prepare_resident(X.h) == pkg.binding.ORBX_ERR_STATE
    X.close()

"""GPU parity of the extraction path (through the C ABI) against the CPU oracle, stage by stage.

Bars (BASELINE.json north_star): pyramid pixels, FAST keypoint sets, quadtree selection bit-exact;
angles within 1e-3 deg (we demand bit-exact); descriptors bit-exact.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [
    # name, rows, cols, nfeatures, lapping, seed
    ("C1-mono", 480, 752, 1000, (0, 1000), 0),
    ("C2-eye", 480, 752, 1200, (0, 0), 3),
    ("C3-fisheye-partial", 512, 512, 1000, (128, 383), 5),
    ("C3-fisheye-full", 512, 512, 1000, (0, 511), 6),
    ("C4-kitti", 376, 1241, 2000, (0, 0), 7),
    ("ini-5x", 480, 752, 5000, (0, 1000), 8),
    ("small", 256, 320, 300, (0, 0), 9),
]


def _cand_arr(c):
    return np.stack([c["x"], c["y"], c["score"]], axis=1).astype(np.int32) if len(c) else np.empty((0, 3), np.int32)


@pytest.mark.parametrize("name,rows,cols,nf,lap,seed", CASES)
def test_extract_stages_match_oracle(pkg, orc, oracle, name, rows, cols, nf, lap, seed):
    img = pkg.synth.synth_frame(rows, cols, seed)
    E = orc.Extractor(nf, oracle=oracle)
    mono_o, kps_o, desc_o = E(img, lap)
    X = pkg.ORBextractor(nf)
    mono_g, kps_g, desc_g = X(img, lap)
    # constructor tables
    assert np.array_equal(X.GetScaleFactors(), E.scale_factors)
    assert np.array_equal(X.GetInverseScaleFactors(), E.inv_scale_factors)
    assert np.array_equal(X.GetScaleSigmaSquares(), E.level_sigma2)
    assert np.array_equal(X.GetInverseScaleSigmaSquares(), E.inv_level_sigma2)
    assert np.array_equal(X.features_per_level(), E.features_per_level)
    for l in range(8):
        assert np.array_equal(X.pyramid_level(l, padded=True), E.level_plane(l)), f"pyramid level {l}"
        assert np.array_equal(X.level_keys(l, 0), _cand_arr(E.level_candidates(l))), f"FAST candidates level {l}"
        sel_g, ang_g = X.level_keys(l, 1)
        sel_o, ang_o = E.level_selected(l)
        assert np.array_equal(sel_g, _cand_arr(sel_o)), f"quadtree selection level {l}"
        assert np.array_equal(ang_g, ang_o), f"angles level {l}: max diff {np.abs(ang_g - ang_o).max() if len(ang_o) else 0}"
        b = E.level_blurred(l)
        if b is not None:
            assert np.array_equal(X.blurred_level(l), b), f"blur level {l}"
    assert mono_g == mono_o
    assert len(kps_g) == len(kps_o)
    for fld in kps_o.dtype.names:
        assert np.array_equal(kps_g[fld], kps_o[fld]), fld
    assert np.array_equal(desc_g, desc_o)
    X.close()


def test_batch_equals_single(pkg, orc, oracle):
    rows, cols, nf = 480, 752, 1000
    frames = np.stack([pkg.synth.synth_frame(rows, cols, 100 + s, bands=(s % 2 == 0)) for s in range(6)])
    X = pkg.ORBextractor(nf)
    mono, nkp, kps, desc = X.extract_batch(frames, (0, 0))
    E = orc.Extractor(nf, oracle=oracle)
    for f in range(len(frames)):
        m, k, d = E(frames[f], (0, 0))
        assert mono[f] == m and nkp[f] == len(k)
        assert np.array_equal(kps[f, :nkp[f]], k)
        assert np.array_equal(desc[f, :nkp[f]], d)
    # a second call on the same handle with fewer frames and a non-contiguous row step
    big = np.zeros((2, rows, cols + 40), np.uint8)
    big[:, :, :cols] = frames[:2]
    m2, k2, d2 = X(big[1, :, :cols], (0, 0))
    m, k, d = E(frames[1], (0, 0))
    assert m2 == m and np.array_equal(k2, k) and np.array_equal(d2, d)
    X.close()


def test_batch_larger_than_device_slots(pkg, orc, oracle):
    """More frames than the handle keeps slots for (512): the slots are reused chunk by chunk.  Host results must be
    right for every frame; nothing is reported resident afterwards (ORBX_ERR_STATE instead of another frame's data)."""
    rows, cols, nf, n = 256, 320, 300, 530
    base = np.stack([pkg.synth.synth_frame(rows, cols, 300 + s, bands=False) for s in range(7)])
    frames = np.ascontiguousarray(base[np.arange(n) % 7])
    X = pkg.ORBextractor(nf)
    mono, nkp, kps, desc = X.extract_batch(frames, (0, 0))
    E = orc.Extractor(nf, oracle=oracle)
    ref = [E(base[i], (0, 0)) for i in range(7)]
    assert min(len(r[1]) for r in ref) > 20
    for f in range(n):
        m, k, d = ref[f % 7]
        assert mono[f] == m and nkp[f] == len(k), f"frame {f}"
        assert np.array_equal(kps[f, :nkp[f]], k) and np.array_equal(desc[f, :nkp[f]], d), f"frame {f}"
    with pytest.raises(pkg.OrbxError) as ei:
        X.pyramid_level(0, frame=0)
    assert ei.value.code == pkg.binding.ORBX_ERR_STATE
    # a batch that fits is resident again, frame f in slot f
    mono, nkp, kps, desc = X.extract_batch(frames[:13], (0, 0))
    for f in (0, 5, 12):
        assert np.array_equal(X.pyramid_level(2, frame=f), E_level(E, base[f % 7], 2))
    X.close()


def E_level(E, img, level):
    E(img, (0, 0))
    return E.level_plane(level)[19:-19, 19:-19]


def test_flat_and_low_texture_frames(pkg, orc, oracle):
    """Empty cells, minThFAST retries and levels with no keypoints at all."""
    rows, cols = 480, 752
    flat = np.full((rows, cols), 77, np.uint8)
    rng = np.random.default_rng(4)
    sparse = flat.copy()
    for _ in range(12):
        y, x = int(rng.integers(40, rows - 40)), int(rng.integers(40, cols - 40))
        sparse[y:y + 9, x:x + 9] = 200
    grad = np.tile(np.linspace(0, 255, cols).astype(np.uint8), (rows, 1))
    X = pkg.ORBextractor(1000)
    E = orc.Extractor(1000, oracle=oracle)
    for img in (flat, sparse, grad):
        mg, kg, dg = X(img, (0, 0))
        mo, ko, do = E(img, (0, 0))
        assert mg == mo and np.array_equal(kg, ko) and np.array_equal(dg, do)
    X.close()


def test_error_conventions(pkg):
    X = pkg.ORBextractor(1000)
    r, k, d = X(np.empty((0, 0), np.uint8))
    assert r == -1 and len(k) == 0          # operator() returns -1 for an empty image (src/ORBextractor.cc:1538)
    with pytest.raises(pkg.OrbxError) as ei:
        X(np.zeros((60, 60), np.uint8))     # top pyramid level narrower than one 35-px cell
    assert ei.value.code == pkg.binding.ORBX_ERR_GEOMETRY
    with pytest.raises(pkg.OrbxError):
        pkg.ORBextractor(1000, nlevels=99)
    X.close()


@pytest.mark.parametrize("name,rows,cols,nf", [("C5-4k", 2160, 3840, 8000), ("1080p", 1080, 1920, 4000), ("4k-12000", 2160, 3840, 12000)])
def test_large_frames_match_oracle(pkg, orc, oracle, name, rows, cols, nf):
    """C5-shaped frames: the level-0 quadtree tables exceed 64 KB of shared memory (own launch of 1024-thread CTAs,
    tables in shared memory up to 220 KB - C5: 160 KB - and in global memory beyond: 12000 features)."""
    img = pkg.synth.synth_frame(rows, cols, 17)
    E = orc.Extractor(nf, oracle=oracle)
    mono_o, kps_o, desc_o = E(img, (0, 0))
    X = pkg.ORBextractor(nf)
    mono_g, kps_g, desc_g = X(img, (0, 0))
    for l in range(8):
        sel_g, ang_g = X.level_keys(l, 1)
        sel_o, ang_o = E.level_selected(l)
        assert np.array_equal(sel_g, _cand_arr(sel_o)), f"quadtree selection level {l}"
        assert np.array_equal(ang_g, ang_o)
    assert mono_g == mono_o and len(kps_g) == len(kps_o) >= nf
    assert np.array_equal(kps_g, kps_o) and np.array_equal(desc_g, desc_o)
    X.close()


def test_ini_extractor_of_c5_40000_features(pkg, orc, oracle):
    """The mono "Ini" extractor is 5 x nFeatures (reference src/Tracking.cc:1113): C5's 8000 features -> 40 000, level-0
    quota 8687 > 4093, which the quadtree's packed sort key used to refuse (ORBX_ERR_GEOMETRY)."""
    img = pkg.synth.synth_frame(2160, 3840, 23)
    nf = 40000
    E = orc.Extractor(nf, oracle=oracle)
    mono_o, kps_o, desc_o = E(img, (0, 1000))
    X = pkg.ORBextractor(nf)
    assert X.features_per_level()[0] > 4093
    mono_g, kps_g, desc_g = X(img, (0, 1000))
    for l in range(8):
        sel_g, ang_g = X.level_keys(l, 1)
        sel_o, ang_o = E.level_selected(l)
        assert np.array_equal(sel_g, _cand_arr(sel_o)), f"quadtree selection level {l}"
    assert mono_g == mono_o and len(kps_g) == len(kps_o) >= nf
    assert np.array_equal(kps_g, kps_o) and np.array_equal(desc_g, desc_o)
    X.close()


@pytest.mark.parametrize("rows,cols,nf", [(600, 5000, 3000), (4500, 2400, 6000)])
def test_frames_wider_or_taller_than_4111_px(pkg, orc, oracle, rows, cols, nf):
    """A level plane beyond 4095 px of keypoint region (5000-px-wide / 4500-px-tall frames): the 12-bit coordinates of the
    32-bit candidate key used to refuse these (ORBX_ERR_GEOMETRY); keys are 64-bit words with 16-bit coordinates now.
    (A tall frame must keep width / height >= 0.5: below that the reference's nIni = round(width / height) is 0 and it
    divides by zero, src/ORBextractor.cc:695-698.)"""
    img = pkg.synth.synth_frame(rows, cols, 29)
    E = orc.Extractor(nf, oracle=oracle)
    mono_o, kps_o, desc_o = E(img, (0, 0))
    X = pkg.ORBextractor(nf)
    mono_g, kps_g, desc_g = X(img, (0, 0))
    for l in range(8):
        assert np.array_equal(X.pyramid_level(l, padded=True), E.level_plane(l)), f"pyramid level {l}"
        assert np.array_equal(X.level_keys(l, 0), _cand_arr(E.level_candidates(l))), f"FAST candidates level {l}"
        sel_g, ang_g = X.level_keys(l, 1)
        sel_o, ang_o = E.level_selected(l)
        assert np.array_equal(sel_g, _cand_arr(sel_o)), f"quadtree selection level {l}"
    assert max(kps_o["x"].max(), kps_o["y"].max()) > 4200
    assert mono_g == mono_o and len(kps_g) == len(kps_o)
    assert np.array_equal(kps_g, kps_o) and np.array_equal(desc_g, desc_o)
    X.close()


@pytest.mark.parametrize("sf,nlevels,ini,mn", [(1.1, 8, 20, 7), (1.25, 6, 20, 7), (1.3, 5, 15, 5), (1.5, 4, 20, 7), (2.0, 3, 25, 10)])
def test_other_scale_factors_and_thresholds(pkg, orc, oracle, sf, nlevels, ini, mn):
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) away from the YAML defaults, batched (so that the
    batch variants of every kernel run): scale factors beyond ~1.25 do not fit the TMA resize tile (host check,
    LevelGeom::pyr_tma_ok) and beyond ~1.33 not the 8-byte tap window either - both fallbacks must give the same bytes."""
    frames = np.stack([pkg.synth.synth_frame(480, 752, 60 + i) for i in range(24)])
    X = pkg.ORBextractor(900, sf, nlevels, ini, mn)
    E = orc.Extractor(900, sf, nlevels, ini, mn, oracle=oracle)
    _, n, k, d = X.extract_batch(frames, (0, 0))
    for f in (0, 7, 23):
        mo, ko, do = E(frames[f], (0, 0))
        assert len(ko) > 300
        assert n[f] == len(ko) and np.array_equal(k[f, :n[f]], ko) and np.array_equal(d[f, :n[f]], do), (sf, f)
        if f == 23:
            for l in range(nlevels):
                assert np.array_equal(X.pyramid_level(l, frame=f), E.level_plane(l)[19:-19, 19:-19]), f"pyramid level {l}"
    X.close()


def test_equal_count_ties_in_the_ordered_phase(pkg, orc, oracle):
    """A regular dot pattern: many quadtree nodes hold the SAME number of candidates when the ordered final phase
    (:909-992) starts, so the split order is decided by the tie-break (count desc, then list position)."""
    rows, cols = 480, 752
    img = np.full((rows, cols), 60, np.uint8)
    for y in range(28, rows - 28, 12):
        for x in range(28, cols - 28, 12):
            img[y:y + 3, x:x + 3] = 200
    for nf in (300, 1000, 2500):
        X = pkg.ORBextractor(nf)
        E = orc.Extractor(nf, oracle=oracle)
        mg, kg, dg = X(img, (0, 0))
        mo, ko, do = E(img, (0, 0))
        assert len(ko) > 100
        assert mg == mo and np.array_equal(kg, ko) and np.array_equal(dg, do), nf
        X.close()


@pytest.mark.parametrize("npatch,patch,nf,seed", [(20, 24, 2000, 1), (6, 48, 1000, 2), (40, 15, 3000, 3)])
def test_clustered_keypoints_deep_quadtree(pkg, orc, oracle, npatch, patch, nf, seed):
    """Corners only inside a few small noisy patches of an otherwise flat frame, fewer than the per-level quota: the
    quadtree has to split down to single keypoints, far below the depth its candidate bins cover (fast rounds ->
    materialise the labels -> relabel per round); selection, order and descriptors must still equal the oracle's.
    (One single patch would not do: the reference stops as soon as a round leaves the node count unchanged.)"""
    rows, cols = 480, 752
    rng = np.random.default_rng(seed)
    img = np.full((rows, cols), 90, np.uint8)
    for _ in range(npatch):
        blk = rng.integers(0, 2, (patch // 3 + 1, patch // 3 + 1)).astype(np.uint8) * 120 + 40
        tex = np.kron(blk, np.ones((3, 3), np.uint8))
        y0, x0 = int(rng.integers(30, rows - 30 - patch)), int(rng.integers(30, cols - 30 - patch))
        img[y0:y0 + patch, x0:x0 + patch] = tex[:patch, :patch]
    X = pkg.ORBextractor(nf)
    E = orc.Extractor(nf, oracle=oracle)
    mg, kg, dg = X(img, (0, 0))
    mo, ko, do = E(img, (0, 0))
    assert 500 < len(ko) < nf
    for l in range(8):
        sel_g, _ = X.level_keys(l, 1)
        sel_o, _ = E.level_selected(l)
        assert np.array_equal(sel_g, _cand_arr(sel_o)), f"quadtree selection level {l}"
    assert mg == mo and np.array_equal(kg, ko) and np.array_equal(dg, do)
    X.close()


def test_single_frame_path_state_changes(pkg, orc, oracle):
    """The one-chunk path caches a CUDA graph of the pipeline per {frames, lapping area, input pitch, buffers}: calls
    that change any of them (other lapping area, strided rows, another image size = re-reserve, a batch call in
    between, a two-frame call) must re-capture and still equal the oracle."""
    X = pkg.ORBextractor(800)
    E = orc.Extractor(800, oracle=oracle)
    a = pkg.synth.synth_frame(480, 752, 61)
    b = pkg.synth.synth_frame(376, 620, 62)
    wide = np.zeros((480, 800), np.uint8); wide[:, :752] = pkg.synth.synth_frame(480, 752, 63)
    seq = [(a, (0, 0)), (a, (200, 400)), (a, (200, 400)), (wide[:, :752], (200, 400)), (b, (0, 0)), (a, (0, 1000)), (b, (100, 300))]
    for k, (img, lap) in enumerate(seq):
        mg, kg, dg = X(img, lap)
        mo, ko, do = E(np.ascontiguousarray(img), lap)
        assert mg == mo and np.array_equal(kg, ko) and np.array_equal(dg, do), f"call {k}"
        if k == 2:          # a batch in between (same size: the buffers stay, the resident frames change)
            fr = np.stack([pkg.synth.synth_frame(480, 752, 70 + i) for i in range(3)])
            mono, nkp, kps, desc = X.extract_batch(fr, (0, 0))
            for i in range(3):
                mo, ko, do = E(fr[i], (0, 0))
                assert mono[i] == mo and np.array_equal(kps[i, :nkp[i]], ko) and np.array_equal(desc[i, :nkp[i]], do)
    X.close()


@pytest.mark.parametrize("cfg", ["C3", "C4"])
def test_fresh_process_per_geometry(cfg):
    """Function attributes (opt-in shared memory) stick to the process: a geometry whose shared-memory need sits just
    under 48 KB only worked when another geometry had raised the limit before.  Each config in its own process."""
    import subprocess, sys, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import importlib, sys, numpy as np\n"
        f"sys.path.insert(0, {root!r})\n"
        "pkg = importlib.import_module('orb-slam3_b200')\n"
        "from oracle import orc_py as orc\n"
        f"c = pkg.synth.CONFIGS[{cfg!r}]\n"
        "fr = np.stack([pkg.synth.synth_frame(c['rows'], c['cols'], 90 + i) for i in range(8)])\n"
        "X = pkg.ORBextractor(c['nfeatures']); E = orc.Extractor(c['nfeatures'])\n"
        "mono, nkp, kps, desc = X.extract_batch(fr, (0, 0))\n"
        "for i in (0, 7):\n"
        "    mo, ko, do = E(fr[i], (0, 0))\n"
        "    assert len(ko) > 100 and np.array_equal(kps[i, :nkp[i]], ko) and np.array_equal(desc[i, :nkp[i]], do)\n"
        "print('ok')\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]


@pytest.mark.parametrize("nhandles,nframes", [(2, 7), (3, 3), (4, 2), (1, 5)])
def test_extract_batch_multi(pkg, orc, oracle, nhandles, nframes):
    """One host batch over several handles (one per device where the box has them, else several on device 0): frame ranges
    of unequal length, fewer frames than handles; results in frame order and equal to the oracle."""
    import ctypes
    ndev = pkg.binding.load().orbx_device_count()
    rows, cols, nf = 376, 620, 600
    frames = np.stack([pkg.synth.synth_frame(rows, cols, 800 + s, bands=(s % 2 == 0)) for s in range(nframes)])
    X = [pkg.ORBextractor(nf, device=d % ndev) for d in range(nhandles)]
    mono, nkp, kps, desc = pkg.extract_batch_multi(X, frames, (0, 0))
    E = orc.Extractor(nf, oracle=oracle)
    for f in range(nframes):
        m, k, d = E(frames[f], (0, 0))
        assert mono[f] == m and nkp[f] == len(k) > 100, f
        assert np.array_equal(kps[f, :nkp[f]], k) and np.array_equal(desc[f, :nkp[f]], d), f
    # each handle keeps ITS range resident: frame 0 of handle 0 is frame 0 of the batch
    E(frames[0], (0, 0))
    assert np.array_equal(X[0].pyramid_level(1, frame=0), E.level_plane(1)[19:-19, 19:-19])
    # the same handle twice is refused (a handle is not re-entrant)
    L = pkg.binding.load()
    hs = (ctypes.c_void_p * 2)(X[0].h, X[0].h)
    rc = L.orbx_extract_batch_multi(hs, 2, nframes, frames.ctypes.data, rows, cols, cols, rows * cols, 0, 0,
                                    kps.ctypes.data, desc.ctypes.data, kps.shape[1], nkp.ctypes.data, mono.ctypes.data)
    assert rc == pkg.binding.ORBX_ERR_ARG
    for x in X:
        x.close()

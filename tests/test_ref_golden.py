"""Golden vectors PRODUCED BY THE REFERENCE'S OWN CODE (tests/golden/ref_*.npz, made by oracle/make_ref_golden.py from the unmodified
src/ORBextractor.cc and src/Frame.cc compiled in the build container).  They travel with the repo, so these comparisons also run where
/root/reference and oracle/_ref do not exist: the C oracle on the CPU, the CUDA path on the GPU box - both against reference bytes."""
import hashlib
import pathlib

import numpy as np
import pytest

GOLD = pathlib.Path(__file__).resolve().parent / "golden"
EXTRACT = ["c1_mono", "c3_fisheye_partial", "c4_kitti", "small"]


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _case(name, synth):
    g = np.load(GOLD / f"ref_extract_{name}.npz")
    img = synth.synth_frame(int(g["rows"]), int(g["cols"]), int(g["seed"]))
    assert _sha(img) == str(g["input_sha"]), "the synthetic generator changed: regenerate with oracle/make_ref_golden.py"
    return g, img, (int(g["lap"][0]), int(g["lap"][1]))


@pytest.mark.parametrize("name", EXTRACT)
def test_oracle_reproduces_reference_output(pkg_synth, orc, oracle, name):
    g, img, lap = _case(name, pkg_synth)
    E = orc.Extractor(int(g["nfeatures"]), oracle=oracle)
    mono, kps, desc = E(img, lap)
    assert mono == int(g["mono"])
    assert np.array_equal(kps, g["kps"]) and np.array_equal(desc, g["desc"])
    for l in range(8):
        assert _sha(E.level_plane(l)[19:-19, 19:-19]) == str(g["plane_sha"][l]), f"pyramid level {l}"
    assert np.array_equal(E.scale_factors, g["scale_factors"])


def _stereo_inputs(synth):
    g = np.load(GOLD / "ref_stereo_c2.npz")
    c = synth.CONFIGS["C2"]
    L, R = synth.synth_stereo_pair(c["rows"], c["cols"], int(g["seed"]))[:2]
    assert _sha(L) == str(g["left_sha"]) and _sha(R) == str(g["right_sha"])
    return g, c, L, R


def test_oracle_reproduces_reference_stereo(pkg_synth, orc, oracle):
    g, c, L, R = _stereo_inputs(pkg_synth)
    EL, ER = orc.Extractor(c["nfeatures"], oracle=oracle), orc.Extractor(c["nfeatures"], oracle=oracle)
    _, kl, dl = EL(L, (0, 0)); _, kr, dr = ER(R, (0, 0))
    assert np.array_equal(kl, g["kps_left"]) and np.array_equal(kr, g["kps_right"])
    assert _sha(dl) == str(g["desc_left_sha"]) and _sha(dr) == str(g["desc_right_sha"])
    u, d, _, _ = orc.stereo_match(EL, ER, kl, dl, kr, dr, float(g["mb"]), float(g["mbf"]))
    assert np.array_equal(u.view(np.uint32), g["uRight"].view(np.uint32)) and np.array_equal(d.view(np.uint32), g["depth"].view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("name", EXTRACT)
def test_gpu_reproduces_reference_output(pkg, name):
    g, img, lap = _case(name, pkg.synth)
    X = pkg.ORBextractor(int(g["nfeatures"]))
    mono, kps, desc = X(img, lap)
    assert mono == int(g["mono"])
    assert np.array_equal(kps, g["kps"]), f"{np.sum(kps != g['kps'])} keypoints differ from the reference's"
    assert np.array_equal(desc, g["desc"])
    for l in range(8):
        assert _sha(X.pyramid_level(l)) == str(g["plane_sha"][l]), f"pyramid level {l}"
    # the same frame inside a batch (batch kernels, TMA resize / blur)
    frames = np.stack([img] * 20)
    _, n, k, d = X.extract_batch(frames, lap)
    for f in (0, 19):
        assert n[f] == len(g["kps"]) and np.array_equal(k[f, :n[f]], g["kps"]) and np.array_equal(d[f, :n[f]], g["desc"])
    X.close()


@pytest.mark.gpu
def test_gpu_reproduces_reference_stereo(pkg):
    g, c, L, R = _stereo_inputs(pkg.synth)
    XL, XR = pkg.ORBextractor(c["nfeatures"]), pkg.ORBextractor(c["nfeatures"])
    _, kl, dl = XL(L, (0, 0)); _, kr, dr = XR(R, (0, 0))
    assert np.array_equal(kl, g["kps_left"]) and np.array_equal(kr, g["kps_right"])
    assert _sha(dl) == str(g["desc_left_sha"]) and _sha(dr) == str(g["desc_right_sha"])
    uR, depth = pkg.ComputeStereoMatches(XL, XR, float(g["mb"]), float(g["mbf"]), frame=0)
    n = len(kl)
    assert np.array_equal(np.asarray(uR).reshape(-1)[:n].view(np.uint32), g["uRight"].view(np.uint32))
    assert np.array_equal(np.asarray(depth).reshape(-1)[:n].view(np.uint32), g["depth"].view(np.uint32))
    XL.close(); XR.close()


# ---- guided searches: vectors produced by the compiled, unmodified src/ORBmatcher.cc (inputs: tests/search_cases.py) ----
import search_cases

SEARCHES = sorted(search_cases.CASES)


def _search_golden(name, mod):
    g = np.load(GOLD / f"ref_search_{name}.npz")
    c = search_cases.CASES[name](mod)
    assert search_cases.digest(c) == str(g["inputs_sha"]), "the input builders changed: regenerate with oracle/make_ref_golden.py"
    return g, c


@pytest.mark.parametrize("name", SEARCHES)
def test_oracle_reproduces_reference_search(orc, oracle, name):
    g, c = _search_golden(name, orc)
    count, arr = search_cases.run_oracle(name, orc, oracle, c)
    assert count == int(g["count"]) > 100 and np.array_equal(arr, g["result"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", SEARCHES)
def test_gpu_reproduces_reference_search(pkg, name):
    g, c = _search_golden(name, pkg)
    count, arr = search_cases.run_gpu(name, pkg, c)
    assert count == int(g["count"]) and np.array_equal(arr, g["result"]), f"{np.sum(arr != g['result'])} entries differ from the reference's"

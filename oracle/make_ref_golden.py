#!/usr/bin/env python
"""oracle/make_ref_golden.py — TEST INFRASTRUCTURE.  Golden vectors produced by the REFERENCE'S OWN CODE run in this container:
the unmodified /root/reference/src/ORBextractor.cc (oracle/_ref/libref_seq.so: canonical quadtree tie-break, see oracle/ref_wrap.cc) and
the unmodified src/Frame.cc (oracle/_ref/libref_frame.so, Frame::ComputeStereoMatches).  Writes tests/golden/ref_*.npz; they travel to
the GPU box, where /root/reference does not exist, so that tests/test_ref_golden.py can compare the C oracle (CPU) and the CUDA path
(GPU) with bytes the reference itself produced.  Regenerate with:  python oracle/make_ref_golden.py   (needs /root/reference)."""
import ctypes as C
import hashlib
import importlib.util
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "oracle"))
import ref_py  # noqa: E402

spec = importlib.util.spec_from_file_location("orbx_synth", ROOT / "orb-slam3_b200" / "synth.py")
synth = importlib.util.module_from_spec(spec); spec.loader.exec_module(synth)

EXTRACT_CASES = [   # name, rows, cols, nfeatures, lapping area, seed
    ("c1_mono", 480, 752, 1000, (0, 1000), 3),
    ("c3_fisheye_partial", 512, 512, 1000, (80, 400), 5),
    ("c4_kitti", 376, 1241, 2000, (0, 0), 7),
    ("small", 256, 320, 300, (0, 0), 9),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sources_sha():
    f = ROOT / "oracle" / "_ref" / "SOURCES.sha256"
    return f.read_text() if f.exists() else ""


def main():
    ref_py.build()
    out = ROOT / "tests" / "golden"
    for name, rows, cols, nf, lap, seed in EXTRACT_CASES:
        img = synth.synth_frame(rows, cols, seed)
        R = ref_py.RefExtractor(nf, flavour="seq")
        mono, kps, desc = R(img, lap)
        planes = [sha(R.level_plane(l)) for l in range(8)]
        np.savez_compressed(out / f"ref_extract_{name}.npz", rows=rows, cols=cols, nfeatures=nf, lap=np.array(lap), seed=seed,
                            input_sha=sha(img), mono=mono, kps=kps, desc=desc, plane_sha=np.array(planes),
                            scale_factors=R.tables(0), reference_sources=sources_sha())
        print(name, mono, len(kps))
    # Frame::ComputeStereoMatches on the reference's own keypoints / descriptors / pyramids (C2 pair)
    c = synth.CONFIGS["C2"]
    L, Rimg = synth.synth_stereo_pair(c["rows"], c["cols"], 11)[:2]
    EL, ER = ref_py.RefExtractor(c["nfeatures"], flavour="seq"), ref_py.RefExtractor(c["nfeatures"], flavour="seq")
    _, kl, dl = EL(L, (0, 0)); _, kr, dr = ER(Rimg, (0, 0))
    lib = C.CDLL(str(ROOT / "oracle" / "_ref" / "libref_frame.so"))
    VP = C.c_void_p
    lib.ref_frame_stereo_matches.argtypes = [VP] * 6 + [C.c_int, VP, VP, C.c_int, VP, VP, C.c_int, VP, C.c_float, C.c_float, VP, VP]
    pl = [np.ascontiguousarray(EL.level_plane(l, padded=True)) for l in range(8)]
    pr = [np.ascontiguousarray(ER.level_plane(l, padded=True)) for l in range(8)]
    rows = np.array([p.shape[0] - 38 for p in pl], np.int32); cols = np.array([p.shape[1] - 38 for p in pl], np.int32)
    ptr = lambda ps: (C.c_void_p * 8)(*[p.ctypes.data for p in ps])
    stp = lambda ps: (C.c_size_t * 8)(*[p.strides[0] for p in ps])
    sf = np.ascontiguousarray(EL.tables(0), np.float32)
    mbf = np.float32(c["mbf"]); mb = np.float32(mbf / np.float32(c["fx"]))
    uR = np.empty(len(kl), np.float32); depth = np.empty(len(kl), np.float32)
    lib.ref_frame_stereo_matches(ptr(pl), ptr(pr), rows.ctypes.data, cols.ctypes.data, stp(pl), stp(pr), 8, sf.ctypes.data,
                                 kl.ctypes.data, len(kl), dl.ctypes.data, kr.ctypes.data, len(kr), dr.ctypes.data,
                                 float(mb), float(mbf), uR.ctypes.data, depth.ctypes.data)
    np.savez_compressed(out / "ref_stereo_c2.npz", seed=11, mb=mb, mbf=mbf, left_sha=sha(L), right_sha=sha(Rimg), kps_left=kl, kps_right=kr,
                        desc_left_sha=sha(dl), desc_right_sha=sha(dr), uRight=uR, depth=depth, reference_sources=sources_sha())
    print("stereo", int((uR >= 0).sum()), "matches of", len(kl))
    # guided searches through the compiled, unmodified src/ORBmatcher.cc (oracle/_ref/libref_match.so): inputs from tests/search_cases.py
    sys.path.insert(0, str(ROOT / "tests"))
    import copy
    import orc_py
    import search_cases
    from test_oracle_ref_match import _Proxy
    orc_py.build()
    oracle = orc_py.Oracle()
    refo = copy.copy(oracle)
    refo.L = _Proxy(oracle.L, C.CDLL(str(ROOT / "oracle" / "_ref" / "libref_match.so")))
    for name, make in search_cases.CASES.items():
        c = make(orc_py)
        count, arr = search_cases.run_oracle(name, orc_py, refo, c)
        assert any(r.startswith("orc_") for r in refo.L.routed)
        np.savez_compressed(out / f"ref_search_{name}.npz", inputs_sha=search_cases.digest(c), count=count, result=arr,
                            routed=np.array(sorted(refo.L.routed)), reference_sources=sources_sha())
        print(name, count)


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""Generate tests/golden/*.npz with ORACLE A (oracle/ref_cv2.py: the reference's algorithm on the REAL
OpenCV primitives of cv2 4.13.0).  Run in the build container (needs cv2); the fixtures are committed
and pin the dependency-free C oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.

    python oracle/make_golden.py
"""
import hashlib
import pathlib
import sys

import cv2
import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))
import _pkg  # noqa: E402
import ref_cv2  # noqa: E402

synth = _pkg.load_synth_only().synth
OUT = ROOT / "tests" / "golden"

# name, rows, cols, nfeatures, lapping, seed
EXTRACT_CASES = [
    ("c1_mono", 480, 752, 1000, (0, 1000), 0),
    ("c3_fisheye_partial", 512, 512, 1000, (128, 383), 5),
    ("c4_kitti", 376, 1241, 2000, (0, 0), 7),
    ("small", 256, 320, 300, (0, 0), 9),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def triplets(lst):
    return np.array(lst, np.int32).reshape(-1, 3)


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    print("cv2", cv2.__version__)
    for name, rows, cols, nf, lap, seed in EXTRACT_CASES:
        img = synth.synth_frame(rows, cols, seed)
        P = ref_cv2.Params(nf)
        R = ref_cv2.extract(img, P, lap)
        d = dict(cv2_version=cv2.__version__, rows=rows, cols=cols, nfeatures=nf, lap=np.array(lap), seed=seed,
                 input_sha=sha(img), mono=R["mono"], kps=R["kps"], desc=R["desc"],
                 scale_factors=P.sf, inv_scale_factors=P.isf, sigma2=P.sig2, inv_sigma2=P.isig2,
                 nper=np.array(P.nper, np.int32), umax=np.array(P.umax, np.int32),
                 plane_sha=np.array([sha(p) for p in R["planes"]]),
                 blur_sha=np.array([sha(b) if b is not None else "" for b in R["blurs"]]),
                 cand_sha=np.array([sha(triplets(c)) for c in R["cands"]]),
                 ncand=np.array([len(c) for c in R["cands"]], np.int32),
                 nsel=np.array([len(s) for s in R["sels"]], np.int32),
                 sel=np.concatenate([triplets(s) for s in R["sels"]]),
                 angles=np.concatenate([np.array(a, np.float32) for a in R["angles"]]),
                 cell_stats=np.array([[s["cells"], s["retry"], s["empty"]] for s in R["stats"]], np.int32))
        np.savez_compressed(OUT / f"extract_{name}.npz", **d)
        print(name, "n", len(R["kps"]), "mono", R["mono"], "cells/retry/empty L0", R["stats"][0])

    # stereo (C2-shaped pair)
    c = synth.CONFIGS["C2"]
    l, r = synth.synth_stereo_pair(c["rows"], c["cols"], 11)
    P = ref_cv2.Params(c["nfeatures"])
    RL, RR = ref_cv2.extract(l, P, (0, 0)), ref_cv2.extract(r, P, (0, 0))
    mb = c["mbf"] / c["fx"]
    uR, depth = ref_cv2.stereo_matches(RL, RR, P, mb, c["mbf"])
    np.savez_compressed(OUT / "stereo_c2.npz", cv2_version=cv2.__version__, seed=11, mb=np.float32(mb), mbf=np.float32(c["mbf"]),
                        left_sha=sha(l), right_sha=sha(r), kps_left_sha=sha(RL["kps"]), kps_right_sha=sha(RR["kps"]),
                        uRight=uR, depth=depth)
    print("stereo matches", int((uR >= 0).sum()), "of", len(uR))

    # OpenCV primitive known-answer vectors
    rng = np.random.default_rng(0)
    ys = rng.integers(-2_000_000, 2_000_000, 4000).astype(np.float32)
    xs = rng.integers(-2_000_000, 2_000_000, 4000).astype(np.float32)
    ys[:8] = [0, 0, 1, -1, 5, -5, 0, 7]; xs[:8] = [0, 1, 0, 0, 5, -5, -3, 7]
    at = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(ys, xs)], np.float32)
    img = synth.synth_frame(120, 160, 3)
    rs = cv2.resize(img, (133, 100), interpolation=cv2.INTER_LINEAR)
    bl = cv2.GaussianBlur(img, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
    bd = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    fast = {}
    for t in (20, 7):
        k = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(img, None)
        fast[t] = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in k], np.int32).reshape(-1, 3)
    q = rng.integers(0, 256, (64, 32), dtype=np.uint8); t = rng.integers(0, 256, (80, 32), dtype=np.uint8)
    t[40] = t[3]; t[79] = t[3]; q[0] = t[3]
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    knn_idx = np.array([[mm[0].trainIdx, mm[1].trainIdx] for mm in m], np.int32)
    knn_dist = np.array([[int(mm[0].distance), int(mm[1].distance)] for mm in m], np.int32)
    np.savez_compressed(OUT / "primitives.npz", cv2_version=cv2.__version__, atan_y=ys, atan_x=xs, atan=at,
                        img=img, resize_133x100=rs, blur=bl, border=bd, fast20=fast[20], fast7=fast[7],
                        knn_q=q, knn_t=t, knn_idx=knn_idx, knn_dist=knn_dist)
    print("primitives ok")


if __name__ == "__main__":
    main()

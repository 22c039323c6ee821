"""Matching legs of the hot path on the GPU box (north_star: "popc matching ... integer-pipe utilisation"):

  C2  stereo pairs/s: extract both eyes (resident frames) + Frame::ComputeStereoMatches, and the match alone
  C3  BFMatcher kNN(k=2) 1000 x 1000 (ComputeStereoFishEyeMatches) and SearchByProjection of 5 k map points
  big kNN 8192 x 8192 (long enough for pipe counters)

Integer work per descriptor pair: 8 XOR + 8 POPC + 7 IADD + compare ~ 24 int ops.
Run: python tools/bench_match.py [--pairs 256] [--quick]       (prints one JSON object)
"""
import argparse
import json
import pathlib
import sys
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import _pkg  # noqa: E402


def wall(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    t = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t) / reps * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--quick", action="store_true", help="few repetitions (for ncu)")
    args = ap.parse_args()
    reps = 2 if args.quick else args.reps
    warm = 1 if args.quick else 3
    pkg = _pkg.load()
    out = {}

    # ---------------- C2: stereo pairs ----------------
    cfg = pkg.synth.CONFIGS["C2"]
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    mbf, mb = cfg["mbf"], cfg["mbf"] / cfg["fx"]
    nd = 8
    pairs = [pkg.synth.synth_stereo_pair(rows, cols, 100 + s) for s in range(nd)]
    B = args.pairs
    L = np.ascontiguousarray(np.stack([pairs[i % nd][0] for i in range(B)]))
    R = np.ascontiguousarray(np.stack([pairs[i % nd][1] for i in range(B)]))
    XL, XR = pkg.ORBextractor(nf), pkg.ORBextractor(nf)
    XL.upload(L); XR.upload(R)
    lib = pkg.binding.load()

    def extract_both():
        XL.run_resident((0, 0)); XR.run_resident((0, 0))

    def match_only():
        pkg.binding.check(lib.orbx_stereo_match_resident(XL.h, XR.h, mb, mbf)); XL.sync()

    def pair_step():
        extract_both()
        pkg.binding.check(lib.orbx_stereo_match_resident(XL.h, XR.h, mb, mbf))
        XL.sync(); XR.sync()

    extract_both(); XL.sync(); XR.sync()
    ms_match = wall(match_only, reps, warm)
    ms_pair = wall(pair_step, reps, warm)
    _, nkpL, _, _ = XL.download()
    _, nkpR, _, _ = XR.download()
    uR, depth = pkg.ComputeStereoMatches(XL, XR, mb, mbf, frame=None)
    nm = float((uR >= 0).sum()) / B
    pair_gates = float((nkpL.astype(np.int64) * nkpR).sum())
    out["C2_stereo"] = {"pairs_per_step": B, "ms_extract_plus_match": ms_pair, "pairs_per_s": B / (ms_pair * 1e-3),
                        "frames_per_s_extract_equiv": 2 * B / (ms_pair * 1e-3),
                        "ms_match_only": ms_match, "match_share": ms_match / ms_pair,
                        "mean_kp_left": float(nkpL.mean()), "mean_kp_right": float(nkpR.mean()), "mean_matches": nm,
                        "gate_pairs_per_s": pair_gates / (ms_match * 1e-3)}
    XL.close(); XR.close()

    # ---------------- kNN2 ----------------
    rng = np.random.default_rng(7)
    for name, n in (("C3_knn2_1000", 1000), ("knn2_8192", 8192)):
        q = rng.integers(0, 256, (n, 32), dtype=np.uint8); t = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        ms = wall(lambda: pkg.knn2_hamming(q, t), reps, warm)
        out[name] = {"nq": n, "nt": n, "ms_per_call_host_in_out": ms, "pairs_per_s": n * n / (ms * 1e-3),
                     "gintops": 24 * n * n / (ms * 1e-3) / 1e9}

    # ---------------- C3: SearchByProjection vs 5 k map points ----------------
    cfg = pkg.synth.CONFIGS["C3"]
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    X = pkg.ORBextractor(nf)
    img = pkg.synth.synth_frame(rows, cols, 5)
    _, kps, desc = X(img, (0, cols - 1))
    sf = X.GetScaleFactors()
    xy = np.stack([kps["x"], kps["y"]], 1)
    mp = pkg.synth.synth_map_points(desc, xy, kps["octave"], 5000, 11, rows, cols)
    bounds = (0.0, float(cols), 0.0, float(rows))
    ms = wall(lambda: pkg.SearchByProjection(kps, desc, None, bounds, sf, mp, 3.0, 0.8), reps, warm)
    nmatch, _ = pkg.SearchByProjection(kps, desc, None, bounds, sf, mp, 3.0, 0.8)
    out["C3_search_by_projection"] = {"map_points": 5000, "keypoints": int(len(kps)), "ms_per_call_host_in_out": ms, "matches": int(nmatch)}
    X.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()

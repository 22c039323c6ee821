"""Matching and frame-preparation legs of the hot path on the GPU box (north_star: "popc matching ...
integer-pipe utilisation"):

  C2  stereo pairs/s: extract both eyes (resident frames) + Frame::ComputeStereoMatches, and the match alone
  C3  BFMatcher kNN(k=2) 1000 x 1000 (ComputeStereoFishEyeMatches) and SearchByProjection of 5 k map points
  big kNN 8192 x 8192 (long enough for pipe counters)
  prepare: cv::remap + cvtColor fused in front of operator() (EuRoC-shaped raw frames + rectification maps)

Integer work per descriptor pair: 8 XOR + 8 POPC + 7 IADD + compare ~ 24 int ops.
Run: python tools/bench_match.py [--pairs 256] [--quick]       (prints one JSON object)
bench.py imports the legs and adds them to its JSON line ("matching", "prepare").
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


def stereo_leg(pkg, B=256, reps=20, warm=3, device=0):
    """C2: synthetic EuRoC-shaped stereo pairs, 1200 features per eye, extraction + ComputeStereoMatches."""
    cfg = pkg.synth.CONFIGS["C2"]
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    mbf, mb = cfg["mbf"], cfg["mbf"] / cfg["fx"]
    nd = 8
    pairs = [pkg.synth.synth_stereo_pair(rows, cols, 100 + s) for s in range(nd)]
    L = np.ascontiguousarray(np.stack([pairs[i % nd][0] for i in range(B)]))
    R = np.ascontiguousarray(np.stack([pairs[i % nd][1] for i in range(B)]))
    XL, XR = pkg.ORBextractor(nf, device=device), pkg.ORBextractor(nf, device=device)
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
    out = {"workload": "C2: synthetic 752x480 stereo pairs, 1200 features per eye, extract both eyes + ComputeStereoMatches (resident frames)",
           "pairs_per_step": B, "ms_extract_plus_match": ms_pair, "pairs_per_s": B / (ms_pair * 1e-3),
           "ms_match_only": ms_match, "match_share": ms_match / ms_pair,
           "mean_kp_left": float(nkpL.mean()), "mean_kp_right": float(nkpR.mean()), "mean_matches": nm}
    XL.close(); XR.close()
    return out


def knn_leg(pkg, n, reps=20, warm=3, device=0):
    rng = np.random.default_rng(7)
    q = rng.integers(0, 256, (n, 32), dtype=np.uint8); t = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    ms = wall(lambda: pkg.knn2_hamming(q, t, device), reps, warm)
    return {"nq": n, "nt": n, "ms_per_call_host_in_out": ms, "pairs_per_s": n * n / (ms * 1e-3),
            "gintops": 24 * n * n / (ms * 1e-3) / 1e9}


def proj_leg(pkg, reps=20, warm=3, device=0):
    """C3: 512x512 fisheye frame (1000 features) vs 5 k synthetic map points, SearchByProjection."""
    cfg = pkg.synth.CONFIGS["C3"]
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    X = pkg.ORBextractor(nf, device=device)
    img = pkg.synth.synth_frame(rows, cols, 5)
    _, kps, desc = X(img, (0, cols - 1))
    sf = X.GetScaleFactors()
    xy = np.stack([kps["x"], kps["y"]], 1)
    mp = pkg.synth.synth_map_points(desc, xy, kps["octave"], 5000, 11, rows, cols)
    bounds = (0.0, float(cols), 0.0, float(rows))
    ms = wall(lambda: pkg.SearchByProjection(kps, desc, None, bounds, sf, mp, 3.0, 0.8, device=device), reps, warm)
    nmatch, _ = pkg.SearchByProjection(kps, desc, None, bounds, sf, mp, 3.0, 0.8, device=device)
    X.close()
    return {"map_points": 5000, "keypoints": int(len(kps)), "ms_per_call_host_in_out": ms, "matches": int(nmatch)}


def prepare_leg(pkg, B=256, passes=2, reps=5, warm=2, device=0, pinned=None):
    """EuRoC-shaped raw 752x480 gray frames + rectification maps: cv::remap on the device in front of
    operator().  Resident: the prepare kernel alone (GB/s over raw-in + gray-out bytes); e2e: raw host frames
    -> keypoints + descriptors on the host through orbx_prepare_extract_batch."""
    cfg = pkg.synth.CONFIGS["C1"]
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    nd = 16
    base = np.stack([pkg.synth.synth_frame(rows, cols, 500 + s) for s in range(nd)])
    BE = B * passes
    frames = np.ascontiguousarray(base[np.arange(BE) % nd])
    if pinned is not None:
        frames = pinned(frames)
    mx, my = pkg.synth.synth_rectify_maps(rows, cols, seed=1)
    X = pkg.ORBextractor(nf, device=device)
    X.reserve(rows, cols, B)
    X.set_rectify_maps(mx, my)
    X.prepare_upload(frames[:B])
    for _ in range(3):
        X.prepare_resident()
    X.sync()
    K = 20
    X.timer_start()
    for _ in range(K):
        X.prepare_resident()
    ms_k = X.timer_stop() / K
    cap = X.max_keypoints(rows, cols)
    kps = np.empty((BE, cap), pkg.KP_DTYPE); desc = np.empty((BE, cap, 32), np.uint8)
    if pinned is not None:
        kps = pinned(kps.view(np.uint8).reshape(BE, cap, 28)).view(pkg.KP_DTYPE).reshape(BE, cap); desc = pinned(desc)
    ms_e2e = wall(lambda: X.prepare_extract_batch(frames, (0, 1000), kps=kps, desc=desc), reps, warm)
    # the same frames already rectified on the host -> plain extract_batch (what the reference's caller does after its CPU remap)
    X.set_rectify_maps(None, None)
    ms_plain = wall(lambda: X.extract_batch(frames, (0, 1000), kps=kps, desc=desc), reps, warm)
    X.close()
    bytes_alg = 2 * rows * cols
    return {"workload": "752x480 raw gray frames + CV_32FC1 rectification maps -> remap(INTER_LINEAR) -> operator()",
            "frames_resident": B, "prepare_kernel_ms": ms_k, "prepare_kernel_gbs": bytes_alg * B / (ms_k * 1e-3) / 1e9,
            "algorithmic_bytes_per_frame": bytes_alg,
            "e2e_frames_per_call": BE, "e2e_frames_per_s_with_prepare": BE / (ms_e2e * 1e-3),
            "e2e_frames_per_s_plain_extract": BE / (ms_plain * 1e-3)}


def latency_leg(pkg, reps=200, warm=20, device=0):
    """What a SLAM front-end sees per frame: ONE 752x480 frame through operator() (host image in, host keypoints +
    descriptors out), and one stereo pair (two extractors on two host threads, src/Frame.cc:132-137, then
    ComputeStereoMatches with mvuRight / mvDepth back on the host)."""
    import threading
    cfg = pkg.synth.CONFIGS["C2"]
    rows, cols = cfg["rows"], cfg["cols"]
    mbf, mb = cfg["mbf"], cfg["mbf"] / cfg["fx"]
    L, R = pkg.synth.synth_stereo_pair(rows, cols, 7)
    X1 = pkg.ORBextractor(1000, device=device)
    ms_mono = wall(lambda: X1(L, (0, 1000)), reps, warm)
    X1.close()
    XL, XR = pkg.ORBextractor(cfg["nfeatures"], device=device), pkg.ORBextractor(cfg["nfeatures"], device=device)

    def pair():
        t = threading.Thread(target=lambda: XR(R, (0, 0)))
        t.start(); XL(L, (0, 0)); t.join()
        pkg.ComputeStereoMatches(XL, XR, mb, mbf, frame=0)
    ms_pair = wall(pair, reps, warm)
    ms_one = wall(lambda: pkg.ExtractStereo(XL, XR, L, R, mb, mbf), reps, warm)
    XL.close(); XR.close()
    return {"mono_frame_ms": ms_mono, "stereo_pair_ms": ms_pair, "stereo_pair_one_call_ms": ms_one,
            "note": "single 752x480 frame through orbx_extract (host in/out); stereo pair = two extractor threads + "
                    "orbx_stereo_match; one call = orbx_extract_stereo"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--quick", action="store_true", help="few repetitions (for ncu)")
    args = ap.parse_args()
    reps = 2 if args.quick else args.reps
    warm = 1 if args.quick else 3
    pkg = _pkg.load()
    out = {"C2_stereo": stereo_leg(pkg, args.pairs, reps, warm),
           "C3_knn2_1000": knn_leg(pkg, 1000, reps, warm),
           "knn2_8192": knn_leg(pkg, 8192, reps, warm),
           "C3_search_by_projection": proj_leg(pkg, reps, warm),
           "prepare": prepare_leg(pkg, 64 if args.quick else 256, 2, 1 if args.quick else 5, 1 if args.quick else 2),
           "latency": latency_leg(pkg, 10 if args.quick else 200, 2 if args.quick else 20)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

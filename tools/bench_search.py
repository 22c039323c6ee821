"""Per-call latency of the guided searches through the C ABI (host arrays in, host results out) beside
the oracle's scalar loops on one host core.  Run on the GPU box: python tools/bench_search.py"""
import json
import pathlib
import sys
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import _pkg  # noqa: E402
import test_gpu_search as T  # noqa: E402


def timeit(fn, reps=50):
    for _ in range(5):
        fn()
    t = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t) / reps * 1e3


def main():
    pkg = _pkg.load(); orc = _pkg.load_oracle(); orc.build(); oracle = orc.Oracle()
    out = {}
    for n in (1500, 8000):
        rng = np.random.default_rng(n)
        rows, cols = (480, 752) if n < 4000 else (2160, 3840)
        bounds = (0.0, float(cols), 0.0, float(rows))
        lk, ld, pool, src = T._frame(pkg, rng, n, rows, cols)
        ck, cd, perm = T._moved(pkg, rng, lk, ld, rows, cols)
        last = dict(desc=ld, u=lk["x"] + 4, v=lk["y"] - 3, invz=np.full(n, 0.1, np.float32), octave=lk["octave"].astype(np.int32),
                    angle=lk["angle"], valid=np.ones(n, np.uint8), obs=np.ones(n, np.uint8))
        g = orc.Grid(oracle, ck, *bounds)
        r = {}
        r["proj_frame_gpu_ms"] = timeit(lambda: pkg.SearchByProjectionFrame(ck, cd, None, bounds, T.SF, last, 15.0, 0.0, False, False))
        r["proj_frame_cpu_ms"] = timeit(lambda: orc.search_by_projection_frame(g, cd, None, T.SF, last, 15.0, 0.0, False, False), 10)
        prev = np.stack([lk["x"], lk["y"]], 1).astype(np.float32)
        r["init_gpu_ms"] = timeit(lambda: pkg.SearchForInitialization(lk, ld, ck, cd, bounds, prev, 100))
        r["init_cpu_ms"] = timeit(lambda: orc.search_for_initialization(oracle, lk, ld, g, cd, prev, 100), 10)
        node1 = (src % 97).astype(np.int32); node2 = node1[perm]
        kf1 = dict(kps=lk, desc=ld, has_mp=np.zeros(n, np.uint8), uRight=None)
        kf2 = dict(kps=ck, desc=cd, has_mp=np.zeros(n, np.uint8), uRight=None)
        kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = T._feature_vector(rng, node1)
        kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = T._feature_vector(rng, node2)
        F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32); ep = np.array([-1e6, 0], np.float32)
        r["tri_gpu_ms"] = timeit(lambda: pkg.SearchForTriangulation(kf1, kf2, T.SF, T.SIGMA2, ep, F))
        r["tri_cpu_ms"] = timeit(lambda: orc.search_for_triangulation(oracle, kf1, kf2, T.SF, T.SIGMA2, ep, F), 10)
        out[f"n{n}"] = {k: round(v, 4) for k, v in r.items()}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

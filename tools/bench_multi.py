"""Single-process multi-GPU end-to-end leg through orbx_extract_batch_multi: ONE pinned host batch, one extractor handle
per visible GPU, one host thread per handle inside the library, results in frame order.

  python tools/bench_multi.py [--workload C1] [--frames-per-gpu 2048] [--steps 5]

Prints one JSON line per device count (1, 2, 4, ... all): frames/s including H2D of the frames and D2H of the results.
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C1")
    ap.add_argument("--frames-per-gpu", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    import torch
    pkg = _pkg.load()
    cfg = pkg.synth.CONFIGS[a.workload]
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    ndev = pkg.binding.load().orbx_device_count()
    base = np.stack([pkg.synth.synth_frame(rows, cols, s) for s in range(16)])
    counts = [n for n in (1, 2, 4, 8) if n <= ndev]
    if ndev not in counts:
        counts.append(ndev)
    for n in counts:
        B = a.frames_per_gpu * n
        X = [pkg.ORBextractor(nf, device=d) for d in range(n)]
        cap = X[0].max_keypoints(rows, cols)
        t_frames = torch.from_numpy(np.ascontiguousarray(base[np.arange(B) % len(base)])).pin_memory()
        t_kps = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
        t_desc = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
        fr = t_frames.numpy(); kps = t_kps.numpy(); desc = t_desc.numpy()
        nkp = np.empty(B, np.int32); mono = np.empty(B, np.int32)
        import ctypes as C
        hs = (C.c_void_p * n)(*[x.h for x in X])
        L = pkg.binding.load()

        def step():
            pkg.binding.check(L.orbx_extract_batch_multi(hs, n, B, fr.ctypes.data, rows, cols, fr.strides[1], fr.strides[0], 0, 1000,
                                                         kps.ctypes.data, desc.ctypes.data, cap, nkp.ctypes.data, mono.ctypes.data))
        for _ in range(2):
            step()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            step()
        dt = time.perf_counter() - t0
        print(json.dumps({"api": "orbx_extract_batch_multi", "workload": a.workload, "devices": n, "frames_per_call": B,
                          "frames_per_s": a.steps * B / dt, "mean_keypoints": float(nkp.mean()),
                          "gb_per_s_over_pcie": a.steps * B * (rows * cols + cap * 60) / dt / 1e9}), flush=True)
        for x in X:
            x.close()
        del t_frames, t_kps, t_desc


if __name__ == "__main__":
    main()

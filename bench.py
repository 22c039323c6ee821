#!/usr/bin/env python3
"""bench.py — ORB extract+describe frames/s on B200 (BASELINE.json metric), one JSON line.

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA, liborbx.so)
  python bench.py --impl reference --gpus N --steps K ...  reference arm: the reference's CPU path
                                                           (oracle/_ref, all host cores), rank 0 only

A step = one pass of the extraction hot path over one batch of synthetic frames.
  value : frames/s with the batch already resident in HBM (orbx_upload once, orbx_run_resident per
          step), CUDA events on the extractor's stream, max over ranks.
  e2e   : the same metric through the C-ABI call a user makes (orbx_extract_batch) with pinned HOST
          buffers: H2D of the frames and D2H of keypoints+descriptors inside the timed region.
  roofline : dominant stage's algorithmic bytes / its CUDA-event time, against MEASURED_PEAKS.json.
  cpu_baseline : the reference's own src/ORBextractor.cc (oracle/_ref, compiled over an OpenCV stand-in) on the
          box's host cores, with the measured factor between its scalar primitives and real OpenCV stated beside it.
"""
from __future__ import annotations

import argparse
import json
import os
import pathlib
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
import _pkg  # noqa: E402

STAGES = ["pyramid", "fast", "quadtree", "assemble", "blur", "describe"]
N_DISTINCT = 32      # distinct synthetic frames, tiled to the batch


def level_dims(rows, cols, isf):
    return [(int(np.rint(np.float32(rows) * s)), int(np.rint(np.float32(cols) * s))) for s in isf]


STAGE_KERNEL = {"pyramid": "pyr_resize_fast_kernel", "fast": "fast_strip_kernel", "blur": "blur_"}     # blur_tma_kernel / blur_kernel
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def profiled_traffic(stage, frames_per_launch):
    """dram__bytes_read.sum + dram__bytes_write.sum of the stage's kernel from the newest committed full-set
    ncu summary that holds that kernel (profiles/*_full_summary.csv), scaled to this launch's frame count."""
    import csv
    for f in sorted((ROOT / "profiles").glob("*_full_summary.csv"), reverse=True):
        rows = list(csv.reader(open(f)))
        hdr = rows[0]
        cols = {}
        for i, h in enumerate(hdr):
            for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                if h.startswith(key):
                    cols[key] = (i, UNIT.get(h[h.index("[") + 1:h.index("]")], None))
        if len(cols) != 2 or any(u is None for _, u in cols.values()):
            continue
        tot = 0.0; frames = None
        steps = max(1, sum(1 for r in rows[1:] if "describe_kernel" in r[1]))   # a capture may hold several steps
        for r in rows[1:]:
            if STAGE_KERNEL[stage] in r[1]:                 # template instances carry "void ...<...>" around the name
                tot += sum(float(r[i]) * u for i, u in cols.values())
                g = [int(x) for x in r[2].strip("()").split(",")]
                frames = max(g[1], g[2]) if stage != "pyramid" else g[2]
        if tot == 0.0 or not frames:
            continue                                    # a capture of other kernels (matching legs)
        return tot / steps * frames_per_launch / frames, f"{f.name} ({frames} frames per launch, scaled to {frames_per_launch})"
    return None, None


def algorithmic_bytes(rows, cols, isf, nkp_mean):
    """SURVEY.md 8(d) primary figure, per frame and per stage (compulsory stage-wise HBM traffic)."""
    px = [h * w for h, w in level_dims(rows, cols, isf)]
    return {
        "pyramid": sum(px[:-1]) + sum(px[1:]),       # read L0..L6, write L1..L7
        "fast": sum(px),                             # read L0..L7
        "blur": 2 * sum(px),                         # read + write L0..L7
        "outputs": int(round(nkp_mean * 60)),        # 28 B keypoint + 32 B descriptor
    }


def make_frames(pkg, rows, cols, batch, seed0=0):
    base = np.stack([pkg.synth.synth_frame(rows, cols, seed0 + s) for s in range(min(N_DISTINCT, batch))])
    reps = (batch + len(base) - 1) // len(base)
    return np.ascontiguousarray(np.tile(base, (reps, 1, 1))[:batch])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            p = [x.strip() for x in l.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _cpu_extractors(nfeatures, cores):
    """One independent CPU extractor per host thread.  Preferred: oracle/_ref/libref_native.so = the reference's
    own src/ORBextractor.cc compiled -O3 with AVX2/FMA (kind "reference"); else the C port (kind "port")."""
    orc = _pkg.load_oracle()
    import ref_py
    if ref_py.available("native"):
        exts = [ref_py.RefExtractor(nfeatures, flavour="native") for _ in range(cores)]
        cap = 4 * nfeatures + 4096
        bufs = [(np.zeros(cap, ref_py.KP_DTYPE), np.zeros((cap, 32), np.uint8)) for _ in range(cores)]
        calls = [lambda img, lap, e=e, b=b: e.extract_raw(img, lap, b[0], b[1]) for e, b in zip(exts, bufs)]
        return calls, "reference", ("oracle/_ref/libref_native.so = the UNMODIFIED reference src/ORBextractor.cc "
                                    "(g++ -O3 -march=x86-64-v3) over oracle/cvshim, whose OpenCV primitives are scalar C")
    orc.build()
    o = orc.Oracle()
    exts = [orc.Extractor(nfeatures, oracle=o) for _ in range(cores)]
    return [lambda img, lap, e=e: e(img, lap) for e in exts], "port", "oracle/liborc.so (C restatement, -O3 -ffp-contract=off)"


def opencv_primitive_check(frame, nfeatures):
    """SURVEY 8(d) cross-check: the pixel primitives under the CPU baseline are scalar C restatements; the real
    reference links OpenCV's SIMD ones.  Time both on the calls one frame makes (7 resizes, 8 borders, one FAST
    per 35-px cell, 8 blurs; one thread) so that the baseline can be quoted as if it ran on OpenCV."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    orc = _pkg.load_oracle(); orc.build(); o = orc.Oracle()
    E = orc.Extractor(nfeatures, oracle=o)
    E(frame, (0, 0))
    isf = E.inv_scale_factors
    rows, cols = frame.shape
    dims = level_dims(rows, cols, isf)
    levels = [np.ascontiguousarray(E.level_plane(l)[19:-19, 19:-19]) for l in range(len(dims))]
    cells = []
    for l, (h, w) in enumerate(dims):            # the cell loop's sub-images, src/ORBextractor.cc:1053-1106
        minb, maxx, maxy = 16, w - 16, h - 16
        W, H = maxx - minb, maxy - minb
        nc, nr = W // 35, H // 35
        wc, hc = -(-W // nc), -(-H // nr)
        for i in range(nr):
            y0 = minb + i * hc; y1 = min(y0 + hc + 6, maxy)
            if y0 >= maxy - 3:
                continue
            for j in range(nc):
                x0 = minb + j * wc; x1 = min(x0 + wc + 6, maxx)
                if x0 >= maxx - 6:
                    continue
                cells.append(np.ascontiguousarray(levels[l][y0:y1, x0:x1]))
    det = cv2.FastFeatureDetector_create(20, True)

    def t_cv2():
        t0 = time.perf_counter()
        for l in range(1, len(dims)):
            cv2.resize(levels[l - 1], (dims[l][1], dims[l][0]), interpolation=cv2.INTER_LINEAR)
        for l in range(len(dims)):
            cv2.copyMakeBorder(levels[l], 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
            cv2.GaussianBlur(levels[l], (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        for c in cells:
            det.detect(c)
        return time.perf_counter() - t0

    # the scalar primitives exactly as the timed library holds them (same object code, same flags)
    import ctypes as C
    import ref_py
    L = ref_py._lib("native") if ref_py.available("native") else o.L
    vp = C.c_void_p
    L.orc_resize_linear_u8.argtypes = [vp, C.c_int, C.c_int, C.c_size_t, vp, C.c_int, C.c_int, C.c_size_t]
    L.orc_border_reflect101.argtypes = [vp, C.c_int, C.c_int, C.c_size_t, C.c_int]
    L.orc_gaussian7x7_s2.argtypes = [vp, C.c_int, C.c_int, C.c_size_t, vp, C.c_size_t]
    L.orc_fast9_16.argtypes = [vp, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, vp, C.c_int]
    L.orc_fast9_16.restype = C.c_int
    scratch = np.empty(rows * cols * 3 + 64, np.int32)
    padded = [np.zeros((h + 38, w + 38), np.uint8) for h, w in dims]
    outs = [np.empty((h, w), np.uint8) for h, w in dims]

    def t_port():
        t0 = time.perf_counter()
        for l in range(1, len(dims)):
            L.orc_resize_linear_u8(levels[l - 1].ctypes.data, dims[l - 1][0], dims[l - 1][1], levels[l - 1].strides[0],
                                   outs[l].ctypes.data, dims[l][0], dims[l][1], outs[l].strides[0])
        for l in range(len(dims)):
            padded[l][19:-19, 19:-19] = levels[l]
            L.orc_border_reflect101(padded[l].ctypes.data, dims[l][0], dims[l][1], padded[l].strides[0], 19)
            L.orc_gaussian7x7_s2(levels[l].ctypes.data, dims[l][0], dims[l][1], levels[l].strides[0], outs[l].ctypes.data, outs[l].strides[0])
        for c in cells:
            L.orc_fast9_16(c.ctypes.data, c.shape[0], c.shape[1], c.strides[0], 20, 1, scratch.ctypes.data, c.size)
        return time.perf_counter() - t0

    tc = min(t_cv2() for _ in range(5))
    tp = min(t_port() for _ in range(5))
    return {"cv2_ms": tc * 1e3, "port_ms": tp * 1e3, "cells": len(cells), "cv2_version": cv2.__version__}


def cpu_oracle_throughput(pkg, rows, cols, nfeatures, lapping, target_seconds, threads=None):
    """Reference CPU path on the host cores: one independent extractor per thread (ctypes releases the GIL)."""
    cores = threads or (os.cpu_count() or 1)
    frames = make_frames(pkg, rows, cols, min(N_DISTINCT, 8))
    calls, kind, what = _cpu_extractors(nfeatures, cores)
    calls[0](frames[0], lapping)
    t1 = min(_timed(lambda: calls[0](frames[k % len(frames)], lapping)) for k in range(3))
    n = int(max(cores, min(8000, target_seconds * cores / max(t1, 1e-4))))
    n = (n + cores - 1) // cores * cores
    per = n // cores

    def work(i):
        for k in range(per):
            calls[i](frames[(i + k) % len(frames)], lapping)

    th = [threading.Thread(target=work, args=(i,)) for i in range(cores)]
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    out = {"value": n / dt, "unit": "frames/s", "cores": cores, "kind": kind,
           "sample": f"{n} frames {cols}x{rows} ({nfeatures} features) over {cores} threads in {dt:.1f} s; {what}; "
                     f"single-thread latency {t1 * 1e3:.1f} ms/frame",
           "single_thread_ms": t1 * 1e3}
    chk = opencv_primitive_check(frames[0], nfeatures)
    if chk:
        # the same run with OpenCV's own primitives in place of the scalar ones (measured per frame, one thread)
        adj_ms = max(t1 * 1e3 - chk["port_ms"] + chk["cv2_ms"], chk["cv2_ms"])
        out["opencv_primitives"] = chk
        out["port_vs_cv2"] = chk["port_ms"] / chk["cv2_ms"]
        out["opencv_adjusted_value"] = out["value"] * (t1 * 1e3) / adj_ms
        out["sample"] += (f"; its scalar pixel primitives take {chk['port_ms']:.1f} ms/frame where cv2 {chk['cv2_version']} "
                          f"(SIMD, 1 thread) takes {chk['cv2_ms']:.1f} ms ({out['port_vs_cv2']:.2f}x slower): with OpenCV's "
                          f"primitives the same run would post about {out['opencv_adjusted_value']:.0f} frames/s")
    return out


def _timed(fn):
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


def run_reference(args, cfg, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    pkg = _pkg.load_synth_only()
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    lap = (0, 1000) if args.workload == "C1" else (0, 0)
    cores = os.cpu_count() or 1
    frames = make_frames(pkg, rows, cols, N_DISTINCT)
    calls, kind, what = _cpu_extractors(nf, cores)
    per = max(1, args.ref_frames_per_core)

    def step():
        def work(i):
            for k in range(per):
                calls[i](frames[(i * per + k) % len(frames)], lap)
        th = [threading.Thread(target=work, args=(i,)) for i in range(cores)]
        for t in th:
            t.start()
        for t in th:
            t.join()

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    nframes = per * cores * args.steps
    v = nframes / dt
    cpu = {"value": v, "unit": "frames/s", "cores": cores, "kind": kind,
           "sample": f"{per * cores} frames/step x {args.steps} steps, one extractor per host thread; {what}"}
    chk = opencv_primitive_check(frames[0], nf)
    if chk:
        t1 = min(_timed(lambda: calls[0](frames[k % len(frames)], lap)) for k in range(3)) * 1e3
        adj = max(t1 - chk["port_ms"] + chk["cv2_ms"], chk["cv2_ms"])
        cpu.update({"opencv_primitives": chk, "port_vs_cv2": chk["port_ms"] / chk["cv2_ms"], "single_thread_ms": t1,
                    "opencv_adjusted_value": v * t1 / adj})
        cpu["sample"] += (f"; scalar pixel primitives {chk['port_ms']:.1f} ms/frame vs cv2 {chk['cv2_ms']:.1f} ms "
                          f"({cpu['port_vs_cv2']:.2f}x): with OpenCV's primitives about {cpu['opencv_adjusted_value']:.0f} frames/s")
    line = {"impl": "reference", "metric": "ORB extract+describe frames/s @752x480,1000kp", "value": v, "unit": "frames/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, cfg), "frames_per_step": per * cores},
            "cpu_baseline": cpu,
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(w, cfg):
    return (f"{w}: synthetic {cfg['cols']}x{cfg['rows']} grayscale frames, {cfg['nfeatures']} features, 8 levels, "
            f"scale 1.2, FAST 20/7, ORBextractor::operator() (extract+describe)")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C1", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--batch", type=int, default=0, help="frames per step per GPU (0 = auto, > L2)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="bounded CPU-baseline sample")
    ap.add_argument("--ref-frames-per-core", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-legs", action="store_true", help="skip the matching / prepare legs")
    ap.add_argument("--e2e-passes", type=int, default=4, help="host frames per e2e call = passes x batch")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    synth = _pkg.load_synth_only()
    cfg = synth.synth.CONFIGS[args.workload]
    if args.impl == "reference":
        run_reference(args, cfg, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
    pkg = _pkg.load()
    rows, cols, nf = cfg["rows"], cfg["cols"], cfg["nfeatures"]
    lap = (0, 1000) if args.workload == "C1" else (0, 0)      # mono passes vLappingArea={0,1000} (src/Frame.cc:361)
    B = args.batch or {"C1": 512, "C2": 512, "C3": 512, "C4": 384, "C5": 24}[args.workload]

    X = pkg.ORBextractor(nf, device=local_rank)
    isf = X.GetInverseScaleFactors()
    frames = make_frames(pkg, rows, cols, B, seed0=1000 * rank)
    # pinned host staging for the e2e leg
    cap = X.max_keypoints(rows, cols)
    # the e2e leg streams E2E_PASSES * B host frames per call through the B device slots (ring pipeline)
    BE = args.e2e_passes * B
    t_frames = torch.from_numpy(np.ascontiguousarray(np.tile(frames, (args.e2e_passes, 1, 1)))).pin_memory()
    t_kps = torch.empty((BE, cap, 28), dtype=torch.uint8).pin_memory()
    t_desc = torch.empty((BE, cap, 32), dtype=torch.uint8).pin_memory()
    h_frames = t_frames.numpy()
    h_kps = t_kps.numpy().view(pkg.KP_DTYPE).reshape(BE, cap)
    h_desc = t_desc.numpy()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- resident leg (value) ----------------
    X.upload(h_frames[:B])
    for _ in range(args.warmup):
        X.run_resident(lap)
    X.sync()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    l0 = X.launch_count()
    w0 = time.perf_counter()
    X.timer_start()
    for _ in range(args.steps):
        X.run_resident(lap)
    ms = X.timer_stop()
    wall_ms = (time.perf_counter() - w0) * 1e3
    launches = X.launch_count() - l0
    barrier()
    clocks = sampler.stop()
    ms = max_over_ranks(ms)
    mono, nkp, _, _ = X.download()
    nkp_mean = float(nkp.mean())
    value = world * args.steps * B / (ms * 1e-3)

    # ---------------- per-stage times (profiling events, outside the timed region) ----------------
    X.set_profiling(True)
    stage = np.zeros(len(STAGES)); nprof = 5
    for _ in range(nprof):
        X.run_resident(lap); X.sync()
        stage += np.array([X.stage_ms(i) for i in range(len(STAGES))])
    stage /= nprof
    X.set_profiling(False)
    ab = algorithmic_bytes(rows, cols, isf, nkp_mean)
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    hbm_stages = {"pyramid": 0, "fast": 1, "blur": 4}
    stage_gbs = {k: ab[k] * B / (stage[i] * 1e-3) / 1e9 for k, i in hbm_stages.items() if stage[i] > 0}
    dom = max(hbm_stages, key=lambda k: stage[hbm_stages[k]])
    whole = sum(ab.values()) * B / (ms / args.steps * 1e-3) / 1e9
    traffic, traffic_src = profiled_traffic(dom, B)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": stage_gbs[dom], "peak": peak_gbs, "unit": "GB/s",
                "frac": stage_gbs[dom] / peak_gbs, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "bytes_per_launch": ab[dom] * B,
                "stage_ms": {k: float(stage[i]) for i, k in enumerate(STAGES)},
                "stage_gbs": stage_gbs,
                "pipeline_achieved_gbs": whole, "pipeline_frac": whole / peak_gbs,
                "algorithmic_bytes_per_frame": ab}

    # ---------------- end-to-end leg (host buffers through the C ABI) ----------------
    nkp_h = np.empty(BE, np.int32); mono_h = np.empty(BE, np.int32)
    lib = pkg.binding.load()

    def e2e_step():
        pkg.binding.check(lib.orbx_extract_batch(X.h, BE, h_frames.ctypes.data, rows, cols, h_frames.strides[1], h_frames.strides[0],
                                                 lap[0], lap[1], h_kps.ctypes.data, h_desc.ctypes.data, cap,
                                                 nkp_h.ctypes.data, mono_h.ctypes.data))
    for _ in range(max(1, min(args.warmup, 3))):
        e2e_step()
    barrier()
    ksteps = max(1, min(args.steps, 10))
    t0 = time.perf_counter()
    X.timer_start()
    for _ in range(ksteps):
        e2e_step()
    e2e_ms = X.timer_stop()
    e2e_wall = (time.perf_counter() - t0) * 1e3
    barrier()
    e2e_ms = max_over_ranks(max(e2e_ms, e2e_wall))
    e2e = {"value": world * ksteps * BE / (e2e_ms * 1e-3), "unit": "frames/s",
           "h2d_bytes_per_step": int(BE * rows * cols), "d2h_bytes_per_step": int(BE * (cap * 60 + 8)),
           "steps": ksteps, "frames_per_call": BE,
           "api": "orbx_extract_batch (pinned host frames -> pinned host keypoints+descriptors; chunked H2D/compute/D2H ring)"}
    assert np.array_equal(nkp_h[:B], nkp) and np.array_equal(nkp_h[-B:], nkp), "e2e and resident legs disagree"

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_oracle_throughput(pkg, rows, cols, nf, lap, args.cpu_seconds)

    # ---------------- matching + frame-preparation legs (rank 0, N=1 only; outside the timed regions) ----------------
    matching = prepare = None
    if rank == 0 and world == 1 and not args.no_extra_legs:
        X.close()
        sys.path.insert(0, str(ROOT / "tools"))
        import bench_match as bm
        pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
        try:
            matching = {"stereo": bm.stereo_leg(pkg, 256, 10, 2, local_rank), "knn2_1000": bm.knn_leg(pkg, 1000, 20, 3, local_rank),
                        "knn2_8192": bm.knn_leg(pkg, 8192, 10, 2, local_rank), "search_by_projection_5k": bm.proj_leg(pkg, 20, 3, local_rank),
                        "latency": bm.latency_leg(pkg, 100, 10, local_rank)}
            prepare = bm.prepare_leg(pkg, 256, 2, 4, 2, local_rank, pinned=pin)
        except Exception as ex:                                  # the headline line must still print
            matching = matching or {"error": repr(ex)}
            prepare = prepare or {"error": repr(ex)}

    if rank == 0:
        line = {"metric": "ORB extract+describe frames/s @752x480,1000kp", "value": value, "unit": "frames/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": workload_name(args.workload, cfg), "frames_per_step_per_gpu": B,
                           "distinct_frames": min(N_DISTINCT, B), "mean_keypoints_per_frame": nkp_mean,
                           "l2": f"inputs larger than L2: {B * rows * cols / 1e6:.0f} MB of frames + "
                                 f"{B * sum(h * w for h, w in level_dims(rows, cols, isf)) * 2 / 1e6:.0f} MB of pyramid/blur planes per step vs 126 MB L2",
                           "partition": "frames sharded across ranks, no collective"},
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
                "cpu_baseline": cpu, "wall_ms_per_step": wall_ms / args.steps, "matching": matching, "prepare": prepare}
        print(json.dumps(line), flush=True)
    X.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Pinned host<->device copy rates with 1, 2, 4, ... N GPUs copying AT THE SAME TIME (one process per GPU, started
together behind a barrier), to name the limiter of the end-to-end leg when frames are sharded over the GPUs of one box.

  python tools/ubench/pcie_concurrent.py [--gpus 1,2,4,8] [--numa]

Per GPU count: aggregate H2D, D2H and bidirectional GB/s (sum over GPUs of bytes / the slowest GPU's time).  --numa binds
each process to the CPUs of its GPU's NUMA node before it allocates the pinned buffers (first-touch places them there).
Prints one JSON line per configuration; a table goes into DESIGN.md.
"""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import time


def gpu_numa_cpus(idx):
    """CPU list of the NUMA node the GPU hangs off (sysfs), or None."""
    try:
        bus = subprocess.run(["nvidia-smi", f"--id={idx}", "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True).stdout.strip().lower()
        if bus.startswith("0000"):
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        txt = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        cpus = []
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        return node, cpus
    except Exception:
        return None


def worker(rank, world, numa, barrier, out):
    info = None
    if numa:
        info = gpu_numa_cpus(rank)
        if info:
            os.sched_setaffinity(0, info[1])
    import torch
    torch.cuda.set_device(rank)
    n = 185 * 1024 * 1024                      # one 512-frame batch of 752x480 frames
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); h.fill_(1)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    h2 = torch.empty(n // 6, dtype=torch.uint8).pin_memory(); h2.fill_(1)       # results: ~ 1/6 of the frame bytes
    d2 = torch.empty(n // 6, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(2):
        d.copy_(h, non_blocking=True); h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    res = {}
    reps = 8
    for name in ("h2d", "d2h", "both"):
        barrier.wait()
        t0 = time.perf_counter()
        for _ in range(reps):
            if name == "h2d":
                d.copy_(h, non_blocking=True)
            elif name == "d2h":
                h.copy_(d, non_blocking=True)
            else:
                with torch.cuda.stream(s1):
                    d.copy_(h, non_blocking=True)
                with torch.cuda.stream(s2):
                    h2.copy_(d2, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        nbytes = reps * (n if name != "both" else n + n // 6)
        res[name] = (nbytes, dt)
        barrier.wait()
    out.put((rank, res, info[0] if info else None))


def run(world, numa):
    ctx = mp.get_context("spawn")
    barrier = ctx.Barrier(world)
    out = ctx.Queue()
    ps = [ctx.Process(target=worker, args=(r, world, numa, barrier, out)) for r in range(world)]
    for p in ps:
        p.start()
    got = [out.get() for _ in ps]
    for p in ps:
        p.join()
    line = {"gpus": world, "numa_bound": bool(numa), "numa_nodes": sorted({g[2] for g in got if g[2] is not None})}
    for name in ("h2d", "d2h", "both"):
        tot = sum(g[1][name][0] for g in got)
        slow = max(g[1][name][1] for g in got)
        line[name + "_gbs_total"] = round(tot / slow / 1e9, 1)
        line[name + "_gbs_per_gpu"] = round(tot / slow / 1e9 / world, 1)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--numa", action="store_true")
    a = ap.parse_args()
    import torch
    have = torch.cuda.device_count()
    for w in [int(x) for x in a.gpus.split(",")]:
        if w <= have:
            run(w, False)
            if a.numa:
                run(w, True)

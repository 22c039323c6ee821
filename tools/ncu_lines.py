#!/usr/bin/env python
"""Stall samples of one kernel of an ncu report (captured with --import-source on from a -lineinfo build) summed per CUDA source line:
   python tools/ncu_lines.py report.ncu-rep <kernel-id (1-based launch index in the report)> <mangled-name prefix> <cubin> [top]
The cubin comes from `cuobjdump -xelf all orb-slam3_b200/liborbx.so`; line numbers from `nvdisasm -g`."""
import collections, csv, io, re, subprocess, sys


def main():
    rep, kid, sym, cubin = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]
    top = int(sys.argv[5]) if len(sys.argv) > 5 else 25
    dis = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(dis) if l.startswith(".text." + sym))
    cur, m = None, {}
    for l in dis[start + 1:]:
        if l.startswith("//---------------------"):
            break
        mm = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if mm:
            cur = (mm.group(1), int(mm.group(2))); continue
        mm = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
        if mm and cur:
            m[int(mm.group(1), 16)] = cur
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f":::{kid}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = rows[1]; si = h.index("# Samples")
    body = [r for r in rows[2:] if len(r) == len(h)]
    addr = []
    for r in body:
        try:
            addr.append(int(r[0], 16))
        except ValueError:
            addr.append(None)
    base = min(a for a in addr if a is not None)
    agg = collections.Counter(); tot = 0.0
    for r, a in zip(body, addr):
        if a is None:
            continue
        try:
            s = float(r[si] or 0)
        except ValueError:
            continue
        agg[m.get(a - base, ("?", 0))] += s; tot += s
    print(rows[0][1], "samples", int(tot))
    cache = {}
    for (f, ln), v in agg.most_common(top):
        if f not in cache:
            try:
                cache[f] = open(f).read().splitlines()
            except OSError:
                cache[f] = []
        src = cache[f][ln - 1].strip()[:110] if 0 < ln <= len(cache[f]) else ""
        print("%5.1f%%  %s:%d  %s" % (100 * v / tot, f.split("/")[-1], ln, src))


if __name__ == "__main__":
    main()

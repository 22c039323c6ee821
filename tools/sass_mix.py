"""Opcode mix of one kernel from an ncu report's source page (needs --import-source / -lineinfo captures):
   python tools/sass_mix.py report.ncu-rep kernel_name [top]"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
    lines = out.splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
    rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
    h = rows[0]; ci = h.index("Instructions Executed"); si = h.index("Source"); sm = h.index("# Samples")
    mix = collections.Counter(); samp = collections.Counter(); tot = 0
    for r in rows[1:]:
        if len(r) <= ci or not r[ci].isdigit():
            continue
        toks = r[si].split()
        op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
        base = ".".join(op.split(".")[:2]) if op.startswith(("VIMNMX", "IDP", "LDG", "LDS", "STS", "STG", "LDGSTS", "ISETP", "PRMT")) else op.split(".")[0]
        n = int(r[ci]); mix[base] += n; tot += n
        samp[base] += int(r[sm]) if r[sm].isdigit() else 0
    print(f"{kern}: {tot} warp instructions")
    for k, v in mix.most_common(top):
        print(f"  {k:16s} {v:12d} {v / tot:6.3f}   samples {samp[k]}")


if __name__ == "__main__":
    main()

"""Key utilisation / stall figures of the first launch in a full-set ncu report: python tools/ncu_stalls.py rep [row]"""
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps', 'launch__occupancy_limit_blocks']


def main():
    rep = sys.argv[1]; row = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u, r = rows[0], rows[1], rows[2 + row]
    print(r[h.index("Kernel Name")][:60], r[h.index("Grid Size")], r[h.index("Block Size")])
    for k in KEYS:
        if k in h:
            print(f"  {k:75s} {r[h.index(k)]} {u[h.index(k)]}")
    st = [(float(r[i]), k) for i, k in enumerate(h) if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and r[i]]
    for v, k in sorted(st, reverse=True)[:8]:
        print(f"  stall {k[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]:28s} {v:.3f}")


if __name__ == "__main__":
    main()

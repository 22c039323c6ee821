#!/bin/sh
# Round-end measurement batch (one B200): parity suite, smoke, bench lines, launch list, full-set ncu capture, latency / matching legs.
# Usage (from the repo root, under gpurun): sh tools/final_measure.sh <tag>
T=${1:-r2s}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q > $O/${T}_pytest.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/${T}_smoke.log 2>&1
python bench.py > $O/${T}_bench.log 2>&1
python bench.py --impl reference > $O/${T}_bench_ref.log 2>&1
for w in C2 C3 C4 C5; do python bench.py --no-cpu-baseline --workload $w > $O/${T}_w_$w.log 2>&1; done
python tools/latency_probe.py 400 > $O/${T}_latency.log 2>&1
python tools/bench_match.py > $O/${T}_match.log 2>&1
python tools/bench_search.py > $O/${T}_search.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/${T}_launches_raw.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${T}_ncu_launch.log 2>&1
# one warm resident step of 128 frames (4 compute streams x 32 frames = 52 launches): skip the first three steps, capture half of the fourth
ncu --set full --import-source on --clock-control none --launch-skip 156 -c 26 -f -o $O/${T}_full \
    python tools/profile_step.py 128 4 > $O/${T}_ncu_full.log 2>&1
tail -2 $O/${T}_pytest.log; cat $O/${T}_smoke.log | tail -1; tail -1 $O/${T}_bench.log | cut -c1-300

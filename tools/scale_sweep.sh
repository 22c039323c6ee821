set -x
mkdir -p gpurun_out
P=29700
for w in C1 C4 C5; do
  for n in 2 4 8; do
    P=$((P+1))
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P bench.py --gpus $n --workload $w --steps 10 --warmup 3 --no-extra-legs > gpurun_out/r2t_scale_${w}_$n.log 2>&1
    tail -1 gpurun_out/r2t_scale_${w}_$n.log | cut -c1-200
  done
done
python tools/bench_multi.py --frames-per-gpu 2048 --steps 4 > gpurun_out/r2t_multi_C1.log 2>&1; cat gpurun_out/r2t_multi_C1.log

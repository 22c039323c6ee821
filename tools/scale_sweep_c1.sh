P=29800
for n in 2 4 8; do
  P=$((P+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P bench.py --gpus $n --workload C1 --steps 10 --warmup 3 --no-extra-legs > gpurun_out/r2w_scale_C1_$n.log 2>&1
  tail -1 gpurun_out/r2w_scale_C1_$n.log | cut -c1-160
done

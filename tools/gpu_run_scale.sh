#!/bin/bash
# scaling check on one 8-GPU box: N = 8, 4, 2, 1 back to back (b200 arm)
mkdir -p gpurun_out
for N in 8 4 2; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) \
   bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "bench n$N rc=$?"; tail -1 gpurun_out/bench_n$N.log | cut -c1-200
done
timeout 400 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.log 2>&1; echo "bench n1 rc=$?"; tail -1 gpurun_out/bench_n1.log | cut -c1-200

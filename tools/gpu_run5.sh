#!/bin/bash
# session c pass 2: bench with pipelined e2e, ops table with plan-cached combine/split, ncu source capture of the streaming kernel
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-1800
timeout 600 python tools/bench_ops.py --chi 4096 --rows permute,combine,split,tdot --no-ref --out gpurun_out/ops_f64_4096_copy.json > gpurun_out/ops_f64_copy.log 2>&1; echo "ops rc=$?"; tail -8 gpurun_out/ops_f64_copy.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"skinny_gemm_kernel" --launch-skip 6 -c 1 -f -o gpurun_out/skinny_full \
   python tools/gemm_probe.py > gpurun_out/ncu_skinny.log 2>&1; echo "ncu skinny rc=$?"
ls -la gpurun_out | head

#!/bin/bash
# profile pass: bench line, ncu launch list of the same command, ncu --set full of the top kernels
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-400
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.log 2>&1; echo "bench ref rc=$?"; tail -1 gpurun_out/bench_reference.log | cut -c1-400
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:grouped_gemm_kernel --launch-skip 8 -c 2 -f -o gpurun_out/gemm_full \
   python tools/gemm_probe.py > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"skinny_gemm_kernel|copy_blocks_kernel" --launch-skip 12 -c 6 -f -o gpurun_out/stream_full \
   python tools/gemm_probe.py > gpurun_out/ncu_stream.log 2>&1; echo "ncu stream rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"bj_gram_kernel|bj_rotate_kernel|bj_apply_kernel" --launch-skip 300 -c 3 -f -o gpurun_out/bj_full \
   python tools/bench_ops.py --chi 4096 --rows svd --no-ref --reps 2 > gpurun_out/ncu_bj.log 2>&1; echo "ncu bj rc=$?"
timeout 900 python tools/bench_ops.py --chi 4096 --out gpurun_out/ops_f64_4096.json > gpurun_out/ops_f64.log 2>&1; echo "ops f64 rc=$?"; tail -11 gpurun_out/ops_f64.log
timeout 900 python tools/bench_ops.py --chi 1024 --dtype complex128 --out gpurun_out/ops_c128_1024.json > gpurun_out/ops_c128.log 2>&1; echo "ops c128 rc=$?"; tail -11 gpurun_out/ops_c128.log
ls -la gpurun_out | head -30

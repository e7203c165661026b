#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_ops.py --chi 4096 --rows blas1 --out gpurun_out/ops_blas1_f64_4096.json > gpurun_out/ops_blas1.log 2>&1; echo "ops rc=$?"; tail -7 gpurun_out/ops_blas1.log
timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows blas1 --out gpurun_out/ops_blas1_c128_1024.json > gpurun_out/ops_blas1c.log 2>&1; echo "ops rc=$?"; tail -8 gpurun_out/ops_blas1c.log
timeout 200 ncu --set full --clock-control none --import-source on -k regex:'block_dot_kernel|block_lincomb_kernel|scale_axis_kernel' -c 6 \
   -o gpurun_out/blas1_full python tools/bench_ops.py --chi 4096 --rows blas1 --no-ref > gpurun_out/ncu_blas1.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/blas1_full.ncu-rep --page raw --csv > gpurun_out/blas1_full_raw.csv 2>/dev/null; ls -la gpurun_out/blas1_full* | head

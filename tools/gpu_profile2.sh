#!/bin/bash
# refresh: all gpu tests, bench line (graph replay), complete ncu launch list of the eager matvec, ops tables
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-300
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.log 2>&1; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-svd --no-graph > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 400 python tools/bench_ops.py --chi 4096 --out gpurun_out/ops_f64_4096.json > gpurun_out/ops_f64.log 2>&1; echo "ops f64 rc=$?"; tail -10 gpurun_out/ops_f64.log
timeout 400 python tools/bench_ops.py --chi 1024 --dtype complex128 --out gpurun_out/ops_c128_1024.json > gpurun_out/ops_c128.log 2>&1; echo "ops c128 rc=$?"; tail -10 gpurun_out/ops_c128.log

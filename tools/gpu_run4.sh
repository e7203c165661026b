#!/bin/bash
# round-1 session c: gpu tests, bench line, per-row ops table (float64 chi=4096, complex128 chi=1024)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-600
timeout 900 python tools/bench_ops.py --chi 4096 --rows permute,combine,split,tdot,svd,tsvd,eigh --out gpurun_out/ops_f64_4096.json > gpurun_out/ops_f64.log 2>&1; echo "ops f64 rc=$?"; tail -12 gpurun_out/ops_f64.log
timeout 600 python tools/bench_ops.py --chi 4096 --rows qr --out gpurun_out/ops_f64_4096_qr.json > gpurun_out/ops_f64_qr.log 2>&1; echo "ops qr rc=$?"; tail -4 gpurun_out/ops_f64_qr.log
timeout 900 python tools/bench_ops.py --chi 1024 --dtype complex128 --out gpurun_out/ops_c128_1024.json > gpurun_out/ops_c128.log 2>&1; echo "ops c128 rc=$?"; tail -12 gpurun_out/ops_c128.log

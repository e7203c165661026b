#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 300 python tools/bench_ops.py --chi 4096 --rows blas1 --no-ref --out gpurun_out/ops_blas1_f64_4096_v2.json > gpurun_out/ops_blas1_v2.log 2>&1; echo "ops rc=$?"; tail -7 gpurun_out/ops_blas1_v2.log
timeout 300 python tools/lanczos_probe.py --chi 4096 --iters 10 --out gpurun_out/lanczos_4096_v2.json > gpurun_out/lanczos_v2.log 2>&1; echo "lanczos rc=$?"; tail -1 gpurun_out/lanczos_v2.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-250

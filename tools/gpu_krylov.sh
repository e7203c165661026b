#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "krylov or block_dot or block_lincomb or scale_axis or live_reference" > gpurun_out/pytest_krylov.log 2>&1; tail -3 gpurun_out/pytest_krylov.log
timeout 600 python tools/bench_ops.py --chi 4096 --rows blas1 --out gpurun_out/ops_blas1_f64_4096.json > gpurun_out/ops_blas1.log 2>&1; echo "ops rc=$?"; tail -9 gpurun_out/ops_blas1.log
timeout 600 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows blas1 --out gpurun_out/ops_blas1_c128_1024.json > gpurun_out/ops_blas1c.log 2>&1; echo "ops rc=$?"; tail -9 gpurun_out/ops_blas1c.log
timeout 900 python tools/lanczos_probe.py --chi 4096 --iters 10 --ref --out gpurun_out/lanczos_4096.json > gpurun_out/lanczos.log 2>&1; echo "lanczos rc=$?"; tail -3 gpurun_out/lanczos.log
timeout 900 python tools/lanczos_probe.py --chi 1024 --iters 10 --ref --out gpurun_out/lanczos_1024.json > gpurun_out/lanczos1024.log 2>&1; echo "lanczos rc=$?"; tail -1 gpurun_out/lanczos1024.log

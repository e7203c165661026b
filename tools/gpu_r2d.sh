#!/bin/bash
# round 2, call d: blocked QR (two-level, DMMA trailing update, 512-thread CTAs) -- full GPU suite + per-row timings
mkdir -p gpurun_out
CYB_QR_PROFILE=1 python tools/qr_probe.py 974 520 1 2>&1 | tail -3
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2d_pytest.log
tail -3 gpurun_out/r2d_pytest.log
python tools/bench_ops.py --chi 4096 --rows qr --ref-torch --out gpurun_out/r2d_ops_f64_4096.json > gpurun_out/r2d_ops.log 2>&1
python tools/bench_ops.py --chi 1024 --rows qr --out gpurun_out/r2d_ops_f64_1024.json >> gpurun_out/r2d_ops.log 2>&1
cat gpurun_out/r2d_ops_f64_4096.json gpurun_out/r2d_ops_f64_1024.json | cut -c1-330

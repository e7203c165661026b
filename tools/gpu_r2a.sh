#!/bin/bash
# round 2, call a: GPU test-suite with the new benchmark-size parity tests + the cuSOLVER / TorchBlockBackend comparator
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2a_pytest.log
python tools/bench_ops.py --chi 4096 --rows tdot,svd,tsvd,qr,eigh --ref-torch --out gpurun_out/r2a_ops_f64_4096.json > gpurun_out/r2a_ops.log 2>&1
python tools/bench_ops.py --chi 1024 --dtype complex128 --rows tdot,svd,qr,eigh --ref-torch --out gpurun_out/r2a_ops_c128_1024.json >> gpurun_out/r2a_ops.log 2>&1
tail -5 gpurun_out/r2a_pytest.log

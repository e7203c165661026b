#!/bin/bash
# round 2, call f: block Jacobi with two sector groups on two streams
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "svd or eigh or golden or live or full_size or bench_parity" 2>&1 | tail -4
python tools/bench_ops.py --chi 4096 --rows svd,tsvd,eigh --out gpurun_out/r2f_ops_f64_4096.json > gpurun_out/r2f_ops.log 2>&1
python tools/bench_ops.py --chi 1024 --dtype complex128 --rows svd,eigh --out gpurun_out/r2f_ops_c128_1024.json >> gpurun_out/r2f_ops.log 2>&1
cat gpurun_out/r2f_ops_f64_4096.json gpurun_out/r2f_ops_c128_1024.json | cut -c1-260

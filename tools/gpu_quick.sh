#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "svd or eigh or trunc or full_size" 2>&1 | tail -3
for v in 10 13 14 15; do CYB_GEMM_VARIANT=$v timeout 200 python tools/gemm_probe.py 2>&1 | tail -1; done
echo "default lib (KS=256, AC=128):"; timeout 300 python tools/bench_ops.py --chi 4096 --rows svd,tsvd,eigh --no-ref --reps 4 2>&1 | grep "^| a"
for v in 128_64 128_128 256_64 512_128; do echo "variant $v:"; CYB_LIB=$PWD/cyten_b200/libcyb_var_$v.so timeout 300 python tools/bench_ops.py --chi 4096 --rows svd,eigh --no-ref --reps 4 2>&1 | grep "^| a"; done

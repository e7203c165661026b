#!/bin/bash
# quick check: all gpu parity tests + matvec probe (streaming-kernel unit sweep) + complex ops table
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for u in 4 8 16; do CYB_SK_UNITS_PER_SLOT=$u timeout 300 python tools/gemm_probe.py 2>&1 | tail -1 | sed "s/^/ups=$u /"; done
timeout 900 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows svd,tsvd,eigh --out gpurun_out/ops_c128_1024.json 2>&1 | tail -6

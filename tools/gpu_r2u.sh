#!/bin/bash
# round 2, call u: programmatic dependent launch of the block-Jacobi round kernels, on / off
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "svd or eigh or factor or sweep or cabi" 2>&1 | tail -3 | tee gpurun_out/r2u_pytest.log
for pdl in 0 1; do
  CYB_BJ_PDL=$pdl timeout 300 python tools/bench_ops.py --chi 4096 --rows svd,eigh --no-ref --out gpurun_out/r2u_ops_pdl$pdl.json > /dev/null 2>&1
  echo "pdl $pdl"; cut -c1-120 gpurun_out/r2u_ops_pdl$pdl.json
done
timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows svd,eigh --no-ref --out gpurun_out/r2u_ops_c.json > /dev/null 2>&1; cut -c1-120 gpurun_out/r2u_ops_c.json
timeout 200 python tools/svd_once.py 2>&1 | tail -2

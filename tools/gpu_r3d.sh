#!/bin/bash
# round 2, call 3d: complex GEMM with contiguous warp tiles + edge stages (build/libcyb_cedge.so, -DCYB_GEMM_C_EDGE) against the interleaved one
mkdir -p gpurun_out
CYB_LIB=build/libcyb_cedge.so timeout 600 python -m pytest tests -m gpu -x -q -k "gemm or compose or tdot or cabi or heff or golden" 2>&1 | tail -3
for lib in cyten_b200/libcyten_b200.so build/libcyb_cedge.so; do
  echo $lib
  CYB_LIB=$lib timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows tdot --no-ref --out gpurun_out/x.json > /dev/null 2>&1; cut -c1-170 gpurun_out/x.json
  CYB_LIB=$lib timeout 600 python tools/bench_ops.py --chi 8192 --dtype complex128 --hubbard --no-ref --rows tdot --reps 3 --out gpurun_out/y.json > /dev/null 2>&1; cut -c1-170 gpurun_out/y.json
  CYB_LIB=$lib timeout 300 python tools/bench_ops.py --chi 4096 --dtype complex128 --rows tdot --no-ref --out gpurun_out/z.json > /dev/null 2>&1; cut -c1-170 gpurun_out/z.json
done

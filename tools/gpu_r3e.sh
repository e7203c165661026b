#!/bin/bash
# round 2, call 3e: after the GEMM clean-up (one tiling scheme: contiguous warp tiles + edge stages, real and complex)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r3e_pytest.log
timeout 300 python bench.py --no-plugin --no-sweep --no-svd --no-lanczos --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['roofline']['launch_ms'], round(d['roofline']['frac'],4))"
timeout 600 python tools/bench_ops.py --chi 8192 --dtype complex128 --hubbard --no-ref --rows tdot --reps 3 --out gpurun_out/r3e_hubbard_tdot.json > /dev/null 2>&1; cut -c1-170 gpurun_out/r3e_hubbard_tdot.json
timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows tdot --no-ref --out gpurun_out/r3e_c1024_tdot.json > /dev/null 2>&1; cut -c1-170 gpurun_out/r3e_c1024_tdot.json

#!/bin/bash
# round 2, call l: BASELINE config 4 at its size (Hubbard U(1)xU(1), chi = 8192, complex128) + ncu launch list and full captures
mkdir -p gpurun_out
python tools/bench_ops.py --chi 8192 --dtype complex128 --hubbard --no-ref --rows tdot,svd,tsvd,qr,eigh --reps 2 --out gpurun_out/r2l_ops_hubbard_c128_8192.json > gpurun_out/r2l_ops.log 2>&1
cut -c1-300 gpurun_out/r2l_ops_hubbard_c128_8192.json
python tools/bench_ops.py --chi 1024 --dtype complex128 --hubbard --rows tdot,svd,qr,eigh --reps 3 --out gpurun_out/r2l_ops_hubbard_c128_1024.json >> gpurun_out/r2l_ops.log 2>&1
cut -c1-300 gpurun_out/r2l_ops_hubbard_c128_1024.json
tail -3 gpurun_out/r2l_ops.log
# the bench command exits 0 without ncu first (r2i), then: launch list of the same command
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2l_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-svd --no-graph --no-plugin --no-lanczos --no-sweep > gpurun_out/r2l_ncu.log 2>&1
# full capture of the two dominant kernels of the step (one launch each)
ncu --set full --clock-control none --import-source on -k regex:grouped_gemm_kernel -s 6 -c 1 -o gpurun_out/r2l_gemm python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-svd --no-graph --no-plugin --no-lanczos --no-sweep > gpurun_out/r2l_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:middle_apply_kernel -s 3 -c 1 -o gpurun_out/r2l_middle python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-svd --no-graph --no-plugin --no-lanczos --no-sweep > gpurun_out/r2l_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep

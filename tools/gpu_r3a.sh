#!/bin/bash
# round 2, call 3a: ncu --set full captures (one launch each) of the factorisation kernels, the grouped GEMM and the fused middle
# pass, after each program has run once without ncu in this round (r2z / r2r); the launch list of the default bench step
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-svd --no-graph --no-plugin --no-lanczos --no-sweep"
N="ncu --set full --clock-control none --import-source on"
timeout 300 $N -k regex:bj_gram_kernel -s 300 -c 1 -o gpurun_out/r3a_bj_gram python tools/svd_once.py > gpurun_out/r3a_1.log 2>&1
timeout 300 $N -k regex:bj_rotate_kernel -s 300 -c 1 -o gpurun_out/r3a_bj_rotate python tools/svd_once.py > gpurun_out/r3a_2.log 2>&1
timeout 300 $N -k regex:bj_apply_kernel -s 300 -c 1 -o gpurun_out/r3a_bj_apply python tools/svd_once.py > gpurun_out/r3a_3.log 2>&1
timeout 300 $N -k regex:jacobi_svd_kernel -c 4 -o gpurun_out/r3a_jacobi_small python tools/svd_once.py > gpurun_out/r3a_4.log 2>&1
timeout 300 $N -k regex:householder_qr_kernel -c 2 -o gpurun_out/r3a_qr python tools/bench_ops.py --chi 4096 --rows qr --no-ref --reps 1 > gpurun_out/r3a_5.log 2>&1
timeout 300 $N -k regex:grouped_gemm_kernel -s 6 -c 1 -o gpurun_out/r3a_gemm $B > gpurun_out/r3a_6.log 2>&1
timeout 300 $N -k regex:middle_apply_kernel -s 3 -c 1 -o gpurun_out/r3a_middle $B > gpurun_out/r3a_7.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r3a_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-svd --no-graph --no-plugin --no-lanczos --no-sweep > gpurun_out/r3a_8.log 2>&1
ls -la gpurun_out/r3a_*.ncu-rep

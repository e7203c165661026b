#!/bin/bash
# round 2, call 3l: ncu --set full of bj_gram / bj_apply after their last change (r3b), and the launch list of one SVD (final code)
mkdir -p gpurun_out
N="ncu --set full --clock-control none --import-source on"
timeout 200 $N -k regex:bj_gram_kernel -s 300 -c 1 -o gpurun_out/r3l_bj_gram python tools/svd_once.py > gpurun_out/r3l_1.log 2>&1
timeout 200 $N -k regex:bj_apply_kernel -s 300 -c 1 -o gpurun_out/r3l_bj_apply python tools/svd_once.py > gpurun_out/r3l_2.log 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:bj_ -s 600 -c 300 --csv --log-file gpurun_out/r3l_svd_launches.csv python tools/svd_once.py > gpurun_out/r3l_3.log 2>&1
ls -la gpurun_out/r3l_*.ncu-rep

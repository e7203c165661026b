#!/bin/bash
# first GPU pass: parity tests, FP64 tensor-pipe ceiling, bench, ncu launch list, ncu full capture of the GEMM
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 120 tools/dmma_peak > gpurun_out/dmma_peak.jsonl 2>&1; cat gpurun_out/dmma_peak.jsonl
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -2 gpurun_out/bench.log
timeout 600 python bench.py --steps 5 --warmup 3 --chi 1024 --no-cpu-baseline > gpurun_out/bench_1024.log 2>&1; tail -1 gpurun_out/bench_1024.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:grouped_gemm_kernel -c 2 -f -o gpurun_out/gemm_full \
   python bench.py --steps 1 --warmup 1 --no-svd --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out

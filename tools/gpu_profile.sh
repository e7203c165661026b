#!/bin/bash
# profile pass: bench line, ncu launch list of the same command, ncu --set full of the top kernels
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-400
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:grouped_gemm_kernel --launch-skip 8 -c 2 -f -o gpurun_out/gemm_full \
   python tools/gemm_probe.py > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"skinny_gemm_kernel|copy_blocks_kernel" --launch-skip 12 -c 6 -f -o gpurun_out/stream_full \
   python tools/gemm_probe.py > gpurun_out/ncu_stream.log 2>&1; echo "ncu stream rc=$?"
ls -la gpurun_out | head -20

#!/bin/bash
# round 2, call b: GPU test-suite with the plugin tests (cyten API -> CytenB200Backend -> cyb:: kernels, the reference's own
# pytest subset through the plugin on the CUDA library) + the default bench line with the e2e_plugin key
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2b_pytest.log
python bench.py > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
tail -5 gpurun_out/r2b_pytest.log
tail -c 1500 gpurun_out/r2b_bench.json
tail -3 gpurun_out/r2b_bench.err

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_full_size.py -m gpu -x -q 2>&1 | tail -15
for sh in 0/8 1/8 2/8 3/8 4/8 5/8 6/8 7/8 0/4 3/4; do SHARD=$sh timeout 300 python tools/gemm_probe.py 2>&1 | tail -1; done

#!/bin/bash
# round 2, call s: no-allocation truncation entry point alone in a fresh process, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r2s_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2s_all.log

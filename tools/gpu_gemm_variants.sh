#!/bin/bash
# times the chi = 4096 effective-Hamiltonian step with alternative real GEMM tilings (libraries built with -DCYB_GEMM_BM= ... under build/)
for lib in cyten_b200/libcyten_b200.so build/libcyb_v*.so; do
  CYB_LIB=$lib timeout 180 python bench.py --no-plugin --no-sweep --no-svd --no-lanczos --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib', d['ms_per_step'], d['roofline']['launch_ms'], round(d['roofline']['frac'],4), d['config'].get('matches_native_arm', d['config'].get('parity')))"
done

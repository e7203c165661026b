#!/bin/bash
# round 2, call h: grouped GEMM with warp-tile edge skipping, A/B
for w in 0 1; do
  CYB_GEMM_WSKIP=$w python bench.py --no-svd --no-plugin --no-cpu-baseline --steps 20 > gpurun_out/r2h_bench_wskip$w.json 2> gpurun_out/r2h.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/r2h_bench_wskip$w.json").read().strip().splitlines()[-1])
print("wskip=$w", d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["launch_ms"])
PY
done
python -m pytest tests -m gpu -x -q -k "gemm or compose or tdot or heff or matvec" 2>&1 | tail -2

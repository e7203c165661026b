#!/bin/bash
# round 2, call i: fused H_eff (GEMM / cyb_middle_apply / GEMM) -- tests, A/B against the unfused chain, default bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "fused or sharding or full_size or sweep" 2>&1 | tail -3
python bench.py --unfused --no-svd --no-plugin --no-cpu-baseline --steps 20 > gpurun_out/r2i_bench_unfused.json 2> gpurun_out/r2i.err
python bench.py > gpurun_out/r2i_bench.json 2>> gpurun_out/r2i.err
python - <<'PY'
import json
for f in ("gpurun_out/r2i_bench_unfused.json", "gpurun_out/r2i_bench.json"):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["gemm_share_of_step"], d["e2e"]["ms_per_step"], d.get("lanczos"), (d.get("e2e_plugin") or {}).get("ms_per_step"), d.get("svd_ms"))
PY
tail -2 gpurun_out/r2i.err

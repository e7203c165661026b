#!/bin/bash
# round 2, call 3k: cached factorisation plans + plan-only guard: full GPU suite, the factorisation rows, the default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r3k_pytest.log
timeout 300 python tools/bench_ops.py --chi 4096 --rows svd,tsvd,qr,eigh --no-ref --out gpurun_out/r3k_ops.json > /dev/null 2>&1; cut -c1-110 gpurun_out/r3k_ops.json
timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows svd,qr,eigh --no-ref --out gpurun_out/r3k_ops_c.json > /dev/null 2>&1; cut -c1-110 gpurun_out/r3k_ops_c.json
python bench.py > gpurun_out/r3k_bench_n1.json 2> gpurun_out/r3k_bench_n1.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3k_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d.get("svd_ms"), d["roofline"]["frac"], (d.get("sweep") or {}).get("ms"), (d.get("e2e_plugin") or {}).get("ms_per_step"), (d.get("lanczos") or {}).get("ms"))
PY
tail -2 gpurun_out/r3k_bench_n1.err | cut -c1-300

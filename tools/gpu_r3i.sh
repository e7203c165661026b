#!/bin/bash
# round 2, call 3i: HEAD after the eigh / middle_apply / plugin changes: full GPU suite, smoke, the default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r3i_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/r3i_bench_n1.json 2> gpurun_out/r3i_bench_n1.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3i_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d.get("svd_ms"), d["roofline"]["frac"], (d.get("sweep") or {}).get("ms"), d.get("e2e_plugin"), (d.get("lanczos") or {}).get("ms"))
PY
tail -2 gpurun_out/r3i_bench_n1.err | cut -c1-300

#!/bin/bash
# round 2, call e: sweep step (BASELINE config 5) on one GPU: parity tests, bench line with sweep + e2e_plugin keys, chi=16384 line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2e_pytest.log
tail -3 gpurun_out/r2e_pytest.log
python bench.py > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2e_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "svd_ms")}, d["roofline"]["frac"], d["sweep"], {k: d["e2e_plugin"].get(k) for k in ("ms_per_step", "matches_native_arm")})
PY
python tools/bench_sweep.py --chi 4096 > gpurun_out/r2e_sweep_4096_n1.json 2>> gpurun_out/r2e_bench.err
python tools/bench_sweep.py --chi 16384 --reps 1 > gpurun_out/r2e_sweep_16384_n1.json 2>> gpurun_out/r2e_bench.err
cat gpurun_out/r2e_sweep_4096_n1.json gpurun_out/r2e_sweep_16384_n1.json
tail -3 gpurun_out/r2e_bench.err

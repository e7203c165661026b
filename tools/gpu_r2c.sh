#!/bin/bash
# round 2, call c: blocked QR with DMMA trailing updates (qr_blocked.cuh) + plugin fixes
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2c_pytest.log
tail -5 gpurun_out/r2c_pytest.log
python tools/bench_ops.py --chi 4096 --rows qr,eigh --out gpurun_out/r2c_ops_f64_4096.json > gpurun_out/r2c_ops.log 2>&1
python tools/bench_ops.py --chi 1024 --dtype complex128 --rows qr --out gpurun_out/r2c_ops_c128_1024.json >> gpurun_out/r2c_ops.log 2>&1
python tools/bench_ops.py --chi 1024 --rows qr --out gpurun_out/r2c_ops_f64_1024.json >> gpurun_out/r2c_ops.log 2>&1
cat gpurun_out/r2c_ops_f64_4096.json gpurun_out/r2c_ops_c128_1024.json gpurun_out/r2c_ops_f64_1024.json | cut -c1-400
python bench.py --no-svd > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2c_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step")}, d["e2e"]["ms_per_step"], d["e2e_plugin"])
PY

#!/bin/bash
# round 2, call m: after the look-ahead gate / middle_apply specialisation / launch skipping
mkdir -p gpurun_out
python tools/qr_probe.py 974 520 1 | tail -1; python tools/qr_probe.py 3500 1800 1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r2m_pytest.log
python tools/bench_ops.py --chi 4096 --rows qr --no-ref --out gpurun_out/r2m_ops_qr.json > /dev/null 2>&1; cut -c1-140 gpurun_out/r2m_ops_qr.json
python bench.py > gpurun_out/r2m_bench.json 2> gpurun_out/r2m.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2m_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["gemm_share_of_step"], d["e2e"]["ms_per_step"], d["lanczos"]["ms"], d["e2e_plugin"]["ms_per_step"], d["svd_ms"], d["sweep"]["ms"], d["gpu_launches"])
PY
tail -2 gpurun_out/r2m.err

#!/bin/bash
# round 2, call 3m: BASELINE config 4 (Hubbard chi = 8192 complex128) and config 5 (bond update at chi = 16384) with the final code
mkdir -p gpurun_out
timeout 600 python tools/bench_ops.py --chi 8192 --dtype complex128 --hubbard --no-ref --rows tdot,svd,qr,eigh --reps 2 --out gpurun_out/r3m_ops_hubbard_c128_8192.json > gpurun_out/r3m_ops.log 2>&1
cut -c1-130 gpurun_out/r3m_ops_hubbard_c128_8192.json
timeout 600 python tools/bench_sweep.py --chi 16384 --reps 1 > gpurun_out/r3m_sweep_16384_n1.json 2> gpurun_out/r3m_sweep.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3m_sweep_16384_n1.json").read().strip().splitlines()[-1])
print("sweep 16384", d["ms"], {k: round(v, 1) for k, v in d["parts_ms"].items()})
PY
tail -2 gpurun_out/r3m_sweep.err | cut -c1-200

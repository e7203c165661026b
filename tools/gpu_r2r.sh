#!/bin/bash
# round 2, call r: GEMM loader with hoisted invariants; rotation kernel v4 (G and Q in registers, rotation warp)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "svd or eigh or factor or gemm or compose or heff or sweep or cabi" 2>&1 | tail -3 | tee gpurun_out/r2r_pytest.log
for mode in 0 1; do
  CYB_BJ_ROTMODE=$mode timeout 300 python tools/bench_ops.py --chi 4096 --rows svd,eigh,tdot --no-ref --out gpurun_out/r2r_ops_m$mode.json > /dev/null 2>&1
  echo "rotmode $mode"; cut -c1-150 gpurun_out/r2r_ops_m$mode.json
done
CYB_BJ_ROTMODE=1 timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows svd,eigh --no-ref --out gpurun_out/r2r_ops_c.json > /dev/null 2>&1; cut -c1-150 gpurun_out/r2r_ops_c.json
timeout 600 python bench.py --no-plugin --no-sweep > gpurun_out/r2r_bench.json 2> gpurun_out/r2r.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2r_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"], d["e2e"]["ms_per_step"], d.get("svd_ms"))
PY
tail -2 gpurun_out/r2r.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:bj_ -s 600 -c 300 --csv --log-file gpurun_out/r2r.csv python tools/svd_once.py > gpurun_out/r2r.log 2>&1
python - <<'PY'
import csv, statistics, collections
rows = list(csv.reader(open("gpurun_out/r2r.csv")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
idx = {h: i for i, h in enumerate(rows[hi])}
d = collections.defaultdict(list)
for r in rows[hi + 1:]:
    if len(r) < len(rows[hi]): continue
    v = float(r[idx["Metric Value"]].replace(",", "")); u = r[idx["Metric Unit"]]
    d[r[idx["Kernel Name"]][:40]].append(v / 1e3 if u.startswith("n") else v)
for k, v in d.items(): print(k, len(v), "median us", statistics.median(v))
PY

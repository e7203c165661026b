#!/bin/bash
# round 2, call 3b: gram with all chunks in flight, rotate sums 8 partials at once, apply overlaps the Q load with its first tile
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "svd or eigh or factor or sweep or cabi" 2>&1 | tail -3 | tee gpurun_out/r3b_pytest.log
timeout 300 python tools/bench_ops.py --chi 4096 --rows svd,eigh --no-ref --out gpurun_out/r3b_ops.json > /dev/null 2>&1; cut -c1-120 gpurun_out/r3b_ops.json
timeout 300 python tools/bench_ops.py --chi 1024 --dtype complex128 --rows svd,eigh --no-ref --out gpurun_out/r3b_ops_c.json > /dev/null 2>&1; cut -c1-120 gpurun_out/r3b_ops_c.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:bj_ -s 600 -c 300 --csv --log-file gpurun_out/r3b.csv python tools/svd_once.py > gpurun_out/r3b.log 2>&1
tail -2 gpurun_out/r3b.log
python - <<'PY'
import csv, statistics, collections
rows = list(csv.reader(open("gpurun_out/r3b.csv")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
idx = {h: i for i, h in enumerate(rows[hi])}
d = collections.defaultdict(list)
for r in rows[hi + 1:]:
    if len(r) < len(rows[hi]): continue
    v = float(r[idx["Metric Value"]].replace(",", "")); u = r[idx["Metric Unit"]]
    d[r[idx["Kernel Name"]][:40]].append(v / 1e3 if u.startswith("n") else v)
for k, v in d.items(): print(k, len(v), "median us", statistics.median(v))
PY

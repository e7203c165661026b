#!/bin/bash
# round 2, call 3c: the default bench line at N = $1 GPUs (what the driver runs), kept under profiles/r2/scale
n=${1:-1}
mkdir -p gpurun_out
if [ $n = 1 ]; then
  python bench.py > gpurun_out/r3c_bench_n1.json 2> gpurun_out/r3c_bench_n1.err
  python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r3c_bench_reference_n1.json 2> gpurun_out/r3c_bench_reference_n1.err
  python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 20 --warmup 5 --no-lanczos > gpurun_out/r3c_bench_n$n.json 2> gpurun_out/r3c_bench_n$n.err
fi
python - <<PY
import json
d = json.loads(open("gpurun_out/r3c_bench_n$n.json").read().strip().splitlines()[-1])
print($n, d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d.get("svd_ms"), d["roofline"]["frac"] if d.get("roofline") else None, d.get("sharded_check"), (d.get("sweep") or {}).get("ms"), (d.get("e2e_plugin") or {}).get("ms_per_step"), (d.get("lanczos") or {}).get("ms"))
PY
tail -2 gpurun_out/r3c_bench_n$n.err | cut -c1-300

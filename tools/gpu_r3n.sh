#!/bin/bash
# round 2, call 3n (N GPUs): the sharded bond update (config 5) with the final code; $1 = N, $2 = list of chi
n=$1; shift
mkdir -p gpurun_out
for chi in "$@"; do
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) tools/bench_sweep.py --chi $chi --reps 2 > gpurun_out/r3n_sweep_${chi}_n$n.json 2> gpurun_out/r3n_sweep_n$n.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r3n_sweep_${chi}_n$n.json").read().strip().splitlines()[-1])
    print("sweep", $chi, $n, round(d["ms"], 1), {k: round(v, 1) for k, v in d["parts_ms"].items()})
except Exception as e:
    print("sweep", $chi, $n, "failed", e)
PY
done
tail -2 gpurun_out/r3n_sweep_n$n.err | cut -c1-200

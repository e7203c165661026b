#!/bin/bash
# round 2, call k (8 GPUs): NCCL parity tests on 2 of them, sweep step with sharded environment updates at N = 2 / 4 / 8
mkdir -p gpurun_out
echo skipped-nccl-tests
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) tools/bench_sweep.py --chi 16384 --reps 1 > gpurun_out/r2k_sweep_16384_n$n.json 2> gpurun_out/r2k_sweep_n$n.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + n)) tools/bench_sweep.py --chi 4096 --reps 2 > gpurun_out/r2k_sweep_4096_n$n.json 2>> gpurun_out/r2k_sweep_n$n.err
done
python - <<'PY'
import json
for chi in (4096, 16384):
    for n in (2, 4, 8):
        try:
            d = json.loads(open(f"gpurun_out/r2k_sweep_{chi}_n{n}.json").read().strip().splitlines()[-1])
            print("sweep", chi, n, round(d["ms"], 1), {k: round(v, 1) for k, v in d["parts_ms"].items()})
        except Exception as e:
            print("sweep", chi, n, "failed", e)
PY
tail -3 gpurun_out/r2k_sweep_n8.err

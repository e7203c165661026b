#!/bin/bash
# round 2, call j (8 GPUs): scaling of the default bench line (fused H_eff matvec + truncated SVD) and of the sweep step
# (BASELINE config 5) at chi = 16384, N = 1 / 2 / 4 / 8
mkdir -p gpurun_out
for n in 1 2 4 8; do
  if [ $n = 1 ]; then
    python bench.py --gpus 1 --steps 20 --warmup 5 --no-plugin --no-cpu-baseline --no-lanczos --no-sweep > gpurun_out/r2j_bench_n$n.json 2> gpurun_out/r2j_bench_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2j_bench_n$n.json 2> gpurun_out/r2j_bench_n$n.err
  fi
done
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) tools/bench_sweep.py --chi 16384 --reps 1 > gpurun_out/r2j_sweep_16384_n$n.json 2> gpurun_out/r2j_sweep_n$n.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + n)) tools/bench_sweep.py --chi 4096 --reps 2 > gpurun_out/r2j_sweep_4096_n$n.json 2>> gpurun_out/r2j_sweep_n$n.err
done
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/r2j_bench_n{n}.json").read().strip().splitlines()[-1])
        print(n, d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["svd_ms"], d["roofline"]["frac"] if d["roofline"] else None)
    except Exception as e:
        print(n, "failed", e)
for chi in (4096, 16384):
    for n in (2, 4, 8):
        try:
            d = json.loads(open(f"gpurun_out/r2j_sweep_{chi}_n{n}.json").read().strip().splitlines()[-1])
            print("sweep", chi, n, d["ms"], d["parts_ms"])
        except Exception as e:
            print("sweep", chi, n, "failed", e)
PY
tail -3 gpurun_out/r2j_sweep_n8.err

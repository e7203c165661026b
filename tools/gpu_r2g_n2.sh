#!/bin/bash
# round 2, call g (2 GPUs): NCCL parity tests (sharded matvec / tdot / truncated SVD / sweep step vs the unsharded GPU result)
mkdir -p gpurun_out
python -m pytest tests/test_sharding.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2g_pytest_nccl.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2g_bench_n2.json 2> gpurun_out/r2g_bench_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2g_bench_ref_n2.json 2>> gpurun_out/r2g_bench_n2.err
python - <<'PY'
import json
for f in ("gpurun_out/r2g_bench_n2.json", "gpurun_out/r2g_bench_ref_n2.json"):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, {k: d.get(k) for k in ("value", "ms_per_step", "svd_ms", "n_gpus")}, d.get("cpu_baseline", {}).get("cores") if d.get("cpu_baseline") else None)
PY
tail -2 gpurun_out/r2g_bench_n2.err

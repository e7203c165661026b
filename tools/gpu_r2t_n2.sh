#!/bin/bash
# round 2, call t (2 GPUs): all_reduce behind the last GEMM (OverlappedMatvec): NCCL parity test, then the bench line with
# the overlap off / on and two choices of the group fractions
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sharding.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2t_pytest_nccl.log
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
    bench.py --gpus 2 --steps 20 --warmup 5 --no-svd --no-lanczos > gpurun_out/r2t_bench_n2_$tag.json 2> gpurun_out/r2t_bench_n2_$tag.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2t_bench_n2_$tag.json").read().strip().splitlines()[-1])
    print("$tag", d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d.get("sharded_check"), d["gpu_launches"])
except Exception as e:
    print("$tag failed", e)
PY
  tail -3 gpurun_out/r2t_bench_n2_$tag.err | cut -c1-300
}
run off CYB_OVERLAP=0
run on6040 CYB_OVERLAP=1
run on503020 CYB_OVERLAP=1 CYB_OVERLAP_FRACTIONS=0.5,0.3,0.2
run on7030 CYB_OVERLAP=1 CYB_OVERLAP_FRACTIONS=0.7,0.3

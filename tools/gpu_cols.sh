#!/bin/bash
mkdir -p gpurun_out
run() {  # N tag env...
  N=$1; tag=$2; shift 2
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+RANDOM%300)) \
     bench.py --gpus $N --steps 10 --warmup 3 --no-svd --no-cpu-baseline > gpurun_out/cols_${tag}.log 2>&1
  echo "$tag rc=$?"; tail -1 gpurun_out/cols_${tag}.log | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('  ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'eager', round(d['config']['cuda_graph']['ms_per_step_eager'],3))
except Exception as e: print('  parse failed', e)"
}
run 8 n8_cols06 CYB_SPLIT_COLUMNS=1 CYB_COL_HEAVY=0.6
run 8 n8_cols035 CYB_SPLIT_COLUMNS=1 CYB_COL_HEAVY=0.35
run 8 n8_nocols CYB_SPLIT_COLUMNS=0
run 4 n4_cols06 CYB_SPLIT_COLUMNS=1 CYB_COL_HEAVY=0.6
run 2 n2_cols06 CYB_SPLIT_COLUMNS=1 CYB_COL_HEAVY=0.6

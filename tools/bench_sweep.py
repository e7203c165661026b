"""BASELINE config 5: one bond update of a two-site DMRG sweep (cyten_b200.workloads.SweepStep) on N GPUs of one box.

    python tools/bench_sweep.py --chi 16384                         # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        tools/bench_sweep.py --chi 16384                            # 8 GPUs

Rank 0 prints one JSON line: device time of the whole step (CUDA events, max over ranks, best of --reps) and the wall
time of each part (a synchronisation after every part, second pass).  N > 1: H_eff matvec sharded along vR with the heavy
sectors cut into index ranges, truncated SVD sharded by sector with the global cut from the all-gathered singular values,
environment updates replicated."""
import argparse
import json
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chi", type=int, default=4096)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--chi-max", type=int, default=None)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from cyten_b200 import sharding as sh
    from cyten_b200.backend import B200AbelianBackend
    from cyten_b200.workloads import build_sweep_step
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    be = B200AbelianBackend(dev)
    pg = sh.attach_process_group(be, dist.group.WORLD if world > 1 else None)
    st, th = build_sweep_step(args.chi, be, chi_max=args.chi_max, fused=True)
    st = sh.shard_sweep_step(st, pg)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t0 = time.perf_counter()
    r = st.update_bond(th)
    barrier()
    first_ms = (time.perf_counter() - t0) * 1e3
    best = float("inf")
    for _ in range(args.reps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = st.update_bond(th)
        r["LP"].data.buf, r["RP"].data.buf
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        best = min(best, float(ms.item()))
    st.sync = barrier
    st.timings = {}
    st.update_bond(th)
    if rank == 0:
        tot = sum(st.timings.values())
        print(json.dumps({"workload": f"DMRG sweep step (update_bond), U(1), chi={args.chi}, D=5, d=2, float64", "n_gpus": world,
                          "ms": best, "ms_first_call": first_ms, "parts_ms": {k: v * 1e3 for k, v in st.timings.items()},
                          "share": {k: v / tot for k, v in st.timings.items()}, "kept": int(r["S"].data.shapes.sum()),
                          "err": r["err"], "sweeps_max": (int(be.last_info.abs().max().item()) if be.last_info is not None and be.last_info.numel() else None),
                          "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

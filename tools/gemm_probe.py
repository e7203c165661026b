"""GPU probe: mean time of the tensor-core GEMM launches of the chi=4096 H_eff matvec (variant via CYB_GEMM_VARIANT)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from cyten_b200.backend import B200AbelianBackend
from cyten_b200.workloads import build_heff
chi = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
be = B200AbelianBackend("cuda:0")
H, theta, _, _ = build_heff(chi, be)
if os.environ.get("SHARD"):     # "r/N": the rank-local part of the sharded matvec (no collective), single process
    from cyten_b200 import sharding as sh
    r, N = (int(x) for x in os.environ["SHARD"].split("/"))
    pg = sh.ProcessGroup(None)
    pg.rank, pg.world = r, N
    SH = sh.ShardedHEffective(H, pg, split_heavy=os.environ.get("SPLIT", "0") == "1",
                              split_columns=os.environ.get("COLS", "1") == "1")
    theta = SH.shard(theta)    # flop-cost LPT (+ 2-D split of the heavy sectors)
    class _Loc:
        matvec = staticmethod(SH.matvec_local)
    H = _Loc()
    print("chains:", len(SH.chains), "items:", len(SH.items))
for _ in range(3):
    y = H.matvec(theta)
ref = y.data.buf.clone()
be.gemm_events, be.gemm_event_flops = [], []
evs = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); y = H.matvec(theta); e1.record(); evs.append((e0, e1))
torch.cuda.synchronize()
graph_ms = None
if os.environ.get("GRAPH"):
    from cyten_b200.graphs import CapturedStep
    ge, be.gemm_events = be.gemm_events, None
    cap = CapturedStep(H.matvec, theta)
    gev = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); cap(); e1.record(); gev.append((e0, e1))
    torch.cuda.synchronize()
    graph_ms = np.mean([a.elapsed_time(b) for a, b in gev])
    be.gemm_events = ge
ms = [a.elapsed_time(b) for a, b in be.gemm_events]
fl = be.gemm_event_flops
fat = [(m, f) for m, f in zip(ms, fl) if f > 0.05 * max(fl)]
thin = [m for m, f in zip(ms, fl) if f <= 0.05 * max(fl)]
g = np.mean([m for m, _ in fat]); f = np.mean([f for _, f in fat])
per = {}
print(f"graph_ms={graph_ms} shard={os.environ.get('SHARD')} variant={os.environ.get('CYB_GEMM_VARIANT', 'default')} fat_gemm_ms={g:.4f} TFLOPs={f / g / 1e9:.2f} skinny_ms={np.mean(thin):.4f} "
      f"step_ms={np.mean([a.elapsed_time(b) for a, b in evs]):.3f} checksum={float(y.data.buf.double().abs().sum()):.6e}")

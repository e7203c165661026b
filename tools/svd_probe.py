"""GPU probe: per-stage timing and per-sector sweep counts of truncated_svd(theta) at a given chi."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from cyten_b200 import tensors as ct
from cyten_b200.backend import B200AbelianBackend
from cyten_b200.workloads import CONFIG_LEGS, mps_tensor, u1_bond_leg, u1_phys_leg

chi = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
be = B200AbelianBackend("cuda:0")
Q, sigma = CONFIG_LEGS[chi]
v, p = u1_bond_leg(chi, Q, sigma), u1_phys_leg()
A, B = mps_tensor(v, p, v, be, seed=7), mps_tensor(v, p, v, be, seed=8)
th = ct.tdot(A, B, [2], [0]); th.data.buf
tens, ncd, cc, cd = ct._decomposition_prepare(th, False)
tens.data.buf
def timed(f, n=2):
    best = 1e9
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(); torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3, r
t, _ = timed(lambda: ct._decomposition_prepare(th, False)[0].data.buf); print(f"combine_legs {t:.2f} ms")
t, (u, s, vh) = timed(lambda: be.svd(tens, ncd)); print(f"backend.svd {t:.2f} ms")
info = be.last_info.cpu().numpy()
shapes = tens.data.shapes
print("sector shapes / sweeps:", [(int(a), int(b), int(i)) for (a, b), i in zip(shapes, info)])
t, _ = timed(lambda: ct.truncated_svd(th, chi_max=chi, svd_min=1e-14)); print(f"truncated_svd total {t:.2f} ms")
t, _ = timed(lambda: ct.qr(th)); print(f"qr total {t:.2f} ms")

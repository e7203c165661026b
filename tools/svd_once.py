"""One truncated SVD of the bench's rank-chi theta (tdot of two MPS tensors) with CYB_BJ_DEBUG output."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from cyten_b200 import tensors as ct
from cyten_b200.backend import B200AbelianBackend
from cyten_b200.workloads import CONFIG_LEGS, mps_tensor, u1_bond_leg, u1_phys_leg
chi = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
be = B200AbelianBackend("cuda:0")
Q, sigma = CONFIG_LEGS[chi]
v, p = u1_bond_leg(chi, Q, sigma), u1_phys_leg()
th = ct.tdot(mps_tensor(v, p, v, be, seed=7), mps_tensor(v, p, v, be, seed=8), [2], [0])
th.data.buf
for _ in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r = ct.truncated_svd(th, chi_max=chi, svd_min=1e-14); e1.record(); torch.cuda.synchronize()
    print("truncated_svd ms", e0.elapsed_time(e1), "kept", int(r[1].data.shapes.sum()))

"""One backend.svd call of the combined theta at chi (for ncu launch lists)."""
import sys
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
tens, ncd, cc, cd = ct._decomposition_prepare(th, False)
tens.data.buf
torch.cuda.synchronize()
be.svd(tens, ncd)
torch.cuda.synchronize()

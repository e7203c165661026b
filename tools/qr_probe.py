"""Times cyb_batched_qr on nmat copies of an m x n float64 matrix: python tools/qr_probe.py m n nmat"""
import ctypes as C, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from cyten_b200 import _lib
m, n, nmat = (int(x) for x in sys.argv[1:4])
k = min(m, n)
rng = np.random.default_rng(0)
pad = lambda x: (x + 15) // 16 * 16
A = np.zeros(nmat * pad(m * n)); Q = np.zeros(nmat * pad(m * k)); R = np.zeros(nmat * pad(k * n))
desc = np.zeros((nmat, 6), np.int64)
for i in range(nmat):
    A[i * pad(m * n): i * pad(m * n) + m * n] = rng.standard_normal(m * n)
    desc[i] = (m, n, i * pad(m * n), i * pad(m * k), i * pad(k * n), 0)
dA, dQ, dR = (torch.from_numpy(x).cuda() for x in (A, Q, R))
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: C.c_void_p(t.data_ptr())
def run():
    _lib.call("cyb_batched_qr", p(dA), p(dQ), p(dR), desc.ctypes.data_as(C.c_void_p), nmat, 0, st)
run(); torch.cuda.synchronize()
best = 1e9
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
q = dQ.cpu().numpy()[:m * k].reshape(m, k); r = dR.cpu().numpy()[:k * n].reshape(k, n)
print(f"m={m} n={n} nmat={nmat}: {best:.3f} ms  recon err {np.abs(q @ r - A[:m*n].reshape(m, n)).max():.2e}")

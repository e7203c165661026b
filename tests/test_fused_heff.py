"""The fused effective Hamiltonian (cyten_b200/fused.py): GEMM, one ``cyb_middle_apply`` pass, GEMM must give
the matvec of the unfused call sequence (5 permute_legs + 4 compose, toycodes/tenpy_toycodes/d_dmrg.py:70-88) -- which the other
test files pin against the reference -- to rounding (the <= 8 terms per element are summed in another order), and the kernel
itself must match the numpy restatement of the C-ABI on random groups."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest
import torch

from cyten_b200.fused import FusedHEffective
from cyten_b200.workloads import build_heff


@pytest.mark.parametrize("chi,dtype", [(128, "float64"), (128, "complex128"), (1024, "float64")])
def test_fused_matvec_equals_unfused(backend, chi, dtype):
    if chi > 128 and backend.device.type != "cuda":
        pytest.skip("large case on the GPU only")
    H, theta, _, _ = build_heff(chi, backend, dtype=dtype)
    want = H.matvec(theta)
    F = FusedHEffective(H)
    got = F.matvec(theta)
    assert np.array_equal(got.data.block_inds, want.data.block_inds)
    assert got.codomain == want.codomain and got.domain == want.domain and got.labels == want.labels
    wb, gb = want.data.to_numpy_blocks(), got.data.to_numpy_blocks()
    big = max(float(np.abs(b).max()) for b in wb)
    for g, w in zip(gb, wb):
        assert g.shape == w.shape and np.allclose(g, w, rtol=0, atol=1e-13 * big)
    again = F.matvec(theta)                       # cached tables: bit-identical
    for g, a in zip(gb, again.data.to_numpy_blocks()):
        assert np.array_equal(g, a)


@pytest.mark.gpu
def test_fused_matvec_at_benchmark_size(cuda_backend):
    H, theta, _, _ = build_heff(4096, cuda_backend)
    want, got = H.matvec(theta), FusedHEffective(H).matvec(theta)
    assert np.array_equal(got.data.block_inds, want.data.block_inds)
    big = float(want.data.buf.abs().max()) if False else max(float(want.data.block(i).abs().max()) for i in range(want.data.num_blocks))
    for i in range(want.data.num_blocks):
        assert torch.allclose(got.data.block(i), want.data.block(i), rtol=0, atol=1e-13 * big)


@pytest.mark.parametrize("kind", [pytest.param("oracle"), pytest.param("cuda", marks=pytest.mark.gpu)])
@pytest.mark.parametrize("dtype", ["float64", "complex128"])
def test_middle_apply_kernel(kind, dtype):
    """cyb_middle_apply on random groups (ragged nb, nin / nout from 0 to the maximum) against a direct einsum."""
    from oracle import cabi_numpy as O
    if kind == "cuda":
        if not torch.cuda.is_available():
            pytest.skip("no CUDA device")
        from cyten_b200 import _lib
        _lib.load()
        call, dev = _lib.call, torch.device("cuda:0")
    else:
        call, dev = O.call, torch.device("cpu")
    rng = np.random.default_rng(3)
    cx = dtype == "complex128"
    npdt = np.complex128 if cx else np.float64
    groups, desc, coefs, work0 = [], [], [], [0]
    in_total = out_total = 0
    spec = [(3, 700, 5, 8), (1, 13, 16, 16), (4, 256, 1, 1), (2, 257, 0, 3), (5, 100, 8, 2)]
    layout = []
    for na, nb, nin, nout in spec:
        ins = []
        for i in range(nin):
            ins.append((in_total, nb))                # row i: na vectors of nb elements, stride nb
            in_total += na * nb + 3                   # ragged gaps
        outs = []
        for o in range(nout):
            outs.append((out_total, nb))
            out_total += na * nb + 5
        groups.append((na, nb, nin, nout, len(desc), sum(c.size for c in coefs)))
        desc.extend(ins + outs)
        c = rng.standard_normal((nout, nin)) + (1j * rng.standard_normal((nout, nin)) if cx else 0)
        coefs.append(np.asarray(c, dtype=npdt).reshape(-1))
        work0.append(work0[-1] + na * ((nb + 255) // 256))
        layout.append((ins, outs, c))
    x = rng.standard_normal(in_total) + (1j * rng.standard_normal(in_total) if cx else 0)
    x = np.asarray(x, dtype=npdt)
    y = np.full(out_total, np.nan, dtype=npdt)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    d_x, d_y = t(x), t(y)
    d_g, d_d, d_w = t(np.asarray(groups, np.int64)), t(np.asarray(desc, np.int64)), t(np.asarray(work0, np.int64))
    d_c = t(np.concatenate(coefs))
    p = lambda z: C.c_void_p(z.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream if kind == "cuda" else 0)
    call("cyb_middle_apply", p(d_x), p(d_y), p(d_g), p(d_d), p(d_c), p(d_w), len(groups), work0[-1], 16, 1 if cx else 0, stream)
    got = d_y.cpu().numpy()
    for (na, nb, nin, nout), (ins, outs, c) in zip(spec, layout):
        X = np.stack([x[off:off + na * nb].reshape(na, nb) for off, _ in ins]) if nin else np.zeros((0, na, nb), npdt)
        want = np.tensordot(c, X, axes=(1, 0)) if nin else np.zeros((nout, na, nb), npdt)
        for o, (off, _) in enumerate(outs):
            assert np.allclose(got[off:off + na * nb].reshape(na, nb), want[o], rtol=1e-13, atol=1e-13)
            assert np.all(np.isnan(got[off + na * nb:off + na * nb + 5]))      # the gaps are never written

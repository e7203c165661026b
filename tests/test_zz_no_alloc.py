"""cyb_truncate_select_static: the truncation cut without any allocation at call time (static __device__ scratch).

The entry point has no counterpart in the reference: it is checked against
the existing cyb_truncate_select through the C-ABI, on two calls in a row (the shared scratch keeps the first call's
contents), and against the size limit CYB_TRUNCATE_STATIC_MAX_N of include/cyten_b200.h.  Collected last on purpose
(``zz``): a failure here must not hide the other GPU tests under ``pytest -x``.
"""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def _max_n():
    txt = (ROOT / "include" / "cyten_b200.h").read_text()
    return int(re.search(r"#define\s+CYB_TRUNCATE_STATIC_MAX_N\s+(\d+)", txt).group(1))


def _select(lib, name, torch, S, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error):
    """One call of `name` through the C-ABI on device copies of S; returns (keep, result) on the host."""
    n = len(S)
    d_s = torch.from_numpy(np.ascontiguousarray(S)).cuda()
    d_keep = torch.full((n,), 7, dtype=torch.uint8, device="cuda")       # poisoned outputs
    d_res = torch.full((4,), -123.0, dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = getattr(lib, name)(C.c_void_p(d_s.data_ptr()), n, int(chi_max), int(chi_min), float(degeneracy_tol), float(trunc_cut),
                            float(svd_min), int(minimize_error), C.c_void_p(d_keep.data_ptr()), C.c_void_p(d_res.data_ptr()), stream)
    assert rc == 0, lib.cyb_last_error().decode()
    torch.cuda.synchronize()
    return d_keep.cpu().numpy().astype(bool), d_res.cpu().numpy()


CASES = [
    # n, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error
    (1, 1, 1, 0.0, 0.0, -1.0, 1),
    (37, 20, 1, 0.0, 0.0, 1e-12, 1),
    (300, -1, 1, 0.0, 1e-3, -1.0, 1),
    (5000, 1200, 4, 1e-6, 1e-8, 1e-10, 1),
    (8192, 4096, 1, 0.0, 0.0, 1e-14, 0),
    (65536, 30000, 1, 0.0, 1e-10, 1e-12, 1),     # == CYB_TRUNCATE_STATIC_MAX_N
]


@pytest.mark.gpu
def test_truncate_select_static_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    from cyten_b200 import _lib
    lib = _lib.load()
    assert _max_n() == 65536
    rng = np.random.default_rng(20261105)
    for rep in range(2):                       # second pass: the scratch holds the leftovers of larger and smaller calls
        for n, chi_max, chi_min, dtol, cut, smin, mini in (CASES if rep == 0 else CASES[::-1]):
            S = np.abs(rng.standard_normal(n)) * np.exp(-6.0 * rng.random(n))
            want_keep, want_res = _select(lib, "cyb_truncate_select", torch, S, chi_max, chi_min, dtol, cut, smin, mini)
            for call in range(2):              # two calls in a row on the same input
                keep, res = _select(lib, "cyb_truncate_select_static", torch, S, chi_max, chi_min, dtol, cut, smin, mini)
                # integer / index work: bit-exact; the sums are formed by the same kernel in the same order
                assert np.array_equal(keep, want_keep), (rep, n, call)
                assert np.array_equal(res, want_res), (rep, n, call)
            assert int(want_res[2]) == int(want_keep.sum()) >= 1


@pytest.mark.gpu
def test_truncate_select_static_rejects_sizes_beyond_the_static_scratch():
    import torch
    from cyten_b200 import _lib
    lib = _lib.load()
    n = _max_n() + 1
    d_s = torch.ones(n, dtype=torch.float64, device="cuda")
    d_keep = torch.zeros(n, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(4, dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.cyb_truncate_select_static(C.c_void_p(d_s.data_ptr()), n, -1, 1, 0.0, 0.0, -1.0, 1, C.c_void_p(d_keep.data_ptr()),
                                        C.c_void_p(d_res.data_ptr()), stream)
    assert rc == 3                             # CYB_ERR_LIMIT
    msg = lib.cyb_last_error().decode()
    assert "CYB_TRUNCATE_STATIC_MAX_N" in msg and str(n) in msg
    # the allocating entry point takes the same input
    rc = lib.cyb_truncate_select(C.c_void_p(d_s.data_ptr()), n, -1, 1, 0.0, 0.0, -1.0, 1, C.c_void_p(d_keep.data_ptr()),
                                 C.c_void_p(d_res.data_ptr()), stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert int(d_res.cpu()[2]) == n


def test_static_entry_point_is_declared_and_mirrored():
    """CPU: header, Python binding and the numpy restatement all know the entry point; the limit is the header's."""
    from cyten_b200 import _lib
    from oracle import cabi_numpy
    assert "cyb_truncate_select_static" in _lib.SIGNATURES
    assert _lib.SIGNATURES["cyb_truncate_select_static"] == _lib.SIGNATURES["cyb_truncate_select"]
    S = np.linspace(1.0, 0.01, 50)
    keep_a, keep_b = np.zeros(50, np.uint8), np.zeros(50, np.uint8)
    res_a, res_b = np.zeros(4), np.zeros(4)
    a = lambda x: x.ctypes.data          # the numpy restatement takes addresses, like the library
    assert cabi_numpy.cyb_truncate_select(a(S), 50, 10, 1, 0.0, 0.0, -1.0, 1, a(keep_a), a(res_a)) == 0
    assert cabi_numpy.cyb_truncate_select_static(a(S), 50, 10, 1, 0.0, 0.0, -1.0, 1, a(keep_b), a(res_b)) == 0
    assert keep_a.sum() == 10 and np.array_equal(keep_a, keep_b) and np.array_equal(res_a, res_b)
    assert cabi_numpy.cyb_truncate_select_static(a(S), _max_n() + 1, 10, 1, 0.0, 0.0, -1.0, 1, a(keep_b), a(res_b)) == 3

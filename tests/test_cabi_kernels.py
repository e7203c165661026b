"""Kernel-level parity: every C-ABI entry point of libcyten_b200.so against the numpy oracle
(oracle/cabi_numpy.py) on the same seeded inputs.  Integer / data-movement results bit-exact;
floating point within the tolerance written at each assert.

The ``lib`` fixture runs the same checks with the oracle as "library" on CPU (self-consistency of the
harness + oracle vs plain numpy) and with the CUDA library on the GPU (``-m gpu``).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import cabi_numpy as O

MAX_RANK = 8


class Lib:
    """Uniform caller: uploads numpy arrays, calls ``name`` and downloads the outputs."""

    def __init__(self, kind):
        self.kind = kind
        if kind == "cuda":
            if not torch.cuda.is_available():
                pytest.skip("no CUDA device")
            from cyten_b200 import _lib
            _lib.load()
            self._call = _lib.call
            self.device = torch.device("cuda:0")
        else:
            self._call = O.call
            self.device = torch.device("cpu")
        self.keep = []

    def up(self, a):
        t = torch.from_numpy(np.ascontiguousarray(a)).to(self.device)
        if self.kind != "cuda":
            t = t.clone()
        self.keep.append(t)
        return t

    def empty(self, n, dtype):
        t = torch.zeros(max(int(n), 1), dtype=dtype, device=self.device)
        self.keep.append(t)
        return t

    @staticmethod
    def p(t):
        return C.c_void_p(0 if t is None else t.data_ptr())

    def stream(self):
        return C.c_void_p(torch.cuda.current_stream().cuda_stream if self.kind == "cuda" else 0)

    def call(self, name, *args):
        self._call(name, *args)
        if self.kind == "cuda":
            torch.cuda.synchronize()


@pytest.fixture(params=[pytest.param("oracle"), pytest.param("cuda", marks=pytest.mark.gpu)])
def lib(request):
    return Lib(request.param)


def _rand_rows(rng, n, r, hi=4):
    return rng.integers(0, hi, size=(n, r)).astype(np.int64)


@pytest.mark.parametrize("n,r", [(0, 3), (1, 2), (7, 1), (50, 3), (300, 4), (1500, 5), (4000, 2)])
def test_lexsort_take_rowdiff_pack(lib, n, r):
    rng = np.random.default_rng(n * 10 + r)
    rows = _rand_rows(rng, n, r)
    d_rows, d_perm = lib.up(rows), lib.empty(n, torch.int64)
    lib.call("cyb_lexsort_rows", lib.p(d_rows), n, r, lib.p(d_perm), lib.stream())
    perm = d_perm.cpu().numpy()[:n]
    assert np.array_equal(perm, np.lexsort(rows.T) if n else np.zeros(0, np.int64))   # stable, last col primary
    if n == 0:
        return
    d_sorted = lib.empty(n * r, torch.int64)
    lib.call("cyb_take_rows", lib.p(d_rows), r, lib.p(d_perm), n, lib.p(None), r, lib.p(d_sorted), lib.stream())
    srt = d_sorted.cpu().numpy()[:n * r].reshape(n, r)
    assert np.array_equal(srt, rows[perm])
    d_st, d_cnt = lib.empty(n + 1, torch.int64), lib.empty(1, torch.int64)
    lib.call("cyb_row_differences", lib.p(d_sorted), n, r, lib.p(d_st), lib.p(d_cnt), lib.stream())
    cnt = int(d_cnt.cpu()[0])
    want = [0] + [i for i in range(1, n) if np.any(srt[i - 1] != srt[i])] + [n]
    assert d_st.cpu().numpy()[:cnt].tolist() == want
    strides = rng.integers(1, 50, size=r).astype(np.int64)
    d_keys = lib.empty(n, torch.int64)
    lib.call("cyb_pack_rows", lib.p(d_rows), n, r, lib.p(lib.up(strides)), lib.p(d_keys), lib.stream())
    assert np.array_equal(d_keys.cpu().numpy()[:n], rows @ strides)


@pytest.mark.parametrize("na,nb", [(0, 5), (1, 1), (20, 30), (1200, 900)])
def test_common_sorted(lib, na, nb):
    rng = np.random.default_rng(na + nb)
    a = np.unique(rng.integers(0, 3 * max(na, nb, 1), size=na)).astype(np.int64)
    b = np.unique(rng.integers(0, 3 * max(na, nb, 1), size=nb)).astype(np.int64)
    d_pairs, d_cnt = lib.empty(2 * min(len(a), len(b)) + 2, torch.int64), lib.empty(1, torch.int64)
    lib.call("cyb_common_sorted", lib.p(lib.up(a)) if len(a) else lib.p(lib.empty(1, torch.int64)), len(a),
             lib.p(lib.up(b)) if len(b) else lib.p(lib.empty(1, torch.int64)), len(b), lib.p(d_pairs), lib.p(d_cnt), lib.stream())
    cnt = int(d_cnt.cpu()[0])
    got = d_pairs.cpu().numpy()[:2 * cnt].reshape(cnt, 2)
    common, ia, ib = np.intersect1d(a, b, return_indices=True)
    assert np.array_equal(got, np.stack([ia, ib], axis=1).reshape(-1, 2))


def test_fuse_charges(lib):
    rng = np.random.default_rng(5)
    nsym, nlegs, n = 2, 3, 40
    modulus = np.asarray([0, 3], dtype=np.int64)
    nsec = [4, 3, 5]
    sectors = [np.stack([rng.integers(-3, 4, size=k), rng.integers(0, 3, size=k)], axis=1).astype(np.int64) for k in nsec]
    bi = np.stack([rng.integers(0, k, size=n) for k in nsec] + [rng.integers(0, 2, size=n)], axis=1).astype(np.int64)
    cols = np.asarray([2, 0, 1], dtype=np.int64)     # leg f lives in column cols[f]
    legs_secs = [sectors[2], sectors[0], sectors[1]]
    flat = np.concatenate(legs_secs, axis=0)
    sptr = np.concatenate([[0], np.cumsum([len(s) for s in legs_secs])])[:-1].astype(np.int64)
    out = lib.empty(n * nsym, torch.int64)
    lib.call("cyb_fuse_charges", lib.p(lib.up(bi)), n, bi.shape[1], lib.p(lib.up(cols)), nlegs, lib.p(lib.up(flat)),
             lib.p(lib.up(sptr)), lib.p(lib.up(modulus)), nsym, lib.p(out), lib.stream())
    want = sum(s[bi[:, c]] for s, c in zip(legs_secs, cols))
    want[:, 1] %= 3
    assert np.array_equal(out.cpu().numpy().reshape(n, nsym), want)


def _jobs(src_off, dst_off, shapes, ss, ds):
    n, r = shapes.shape
    J = np.zeros((n, 3 + 3 * MAX_RANK), dtype=np.int64)
    J[:, 0], J[:, 1], J[:, 2] = src_off, dst_off, r
    J[:, 3:3 + r] = shapes
    J[:, 3 + MAX_RANK:3 + MAX_RANK + r] = ss
    J[:, 3 + 2 * MAX_RANK:3 + 2 * MAX_RANK + r] = ds
    return J


def hp(a):
    """host pointer of a contiguous numpy array"""
    return a.ctypes.data_as(C.c_void_p)


def _cstr(shape):
    st = np.ones(len(shape), dtype=np.int64)
    for a in range(len(shape) - 2, -1, -1):
        st[a] = st[a + 1] * shape[a + 1]
    return st


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
@pytest.mark.parametrize("shape,perm", [
    ((5,), (0,)), ((7, 9), (1, 0)), ((70, 45), (1, 0)), ((33, 1, 64), (2, 1, 0)), ((3, 40, 2, 50), (1, 3, 0, 2)),
    ((6, 5, 4, 3, 2), (4, 2, 0, 3, 1)), ((130, 3, 100), (0, 2, 1)), ((2, 300), (0, 1)), ((1, 1, 1), (2, 0, 1)),
    ((64, 64), (1, 0)), ((257, 129), (1, 0)), ((4, 1000), (1, 0)), ((3, 5000), (0, 1)), ((5, 3, 700), (1, 0, 2)),
    ((2, 40, 3, 70), (2, 1, 0, 3)), ((300, 310), (1, 0)), ((2, 3, 4, 2, 3, 2, 3, 2), (7, 0, 6, 1, 5, 2, 4, 3)),
])
def test_copy_blocks_permute(lib, dtype, shape, perm):
    """Batched permutes: three blocks of the same rank but different extents in one launch."""
    rng = np.random.default_rng(len(shape) * 7 + shape[0])
    variants = [tuple(shape), tuple(max(1, s - 1) for s in shape), tuple(s + (1 if i % 2 else 0) for i, s in enumerate(shape))]
    blocks = [(rng.standard_normal(s) + (1j * rng.standard_normal(s) if dtype == np.complex128 else 0)).astype(dtype) for s in variants]
    sizes = [b.size for b in blocks]
    offs = np.concatenate([[0], np.cumsum([(s + 15) // 16 * 16 for s in sizes])])
    src = np.zeros(offs[-1], dtype=dtype)
    for b, o in zip(blocks, offs):
        src[o:o + b.size] = b.reshape(-1)
    new_shapes = np.asarray([[s[p] for p in perm] for s in variants], dtype=np.int64)
    ss = np.asarray([_cstr(s)[list(perm)] for s in variants], dtype=np.int64)
    ds = np.asarray([_cstr(ns) for ns in new_shapes], dtype=np.int64)
    J = _jobs(offs[:-1], offs[:-1], new_shapes, ss, ds)
    d_src, d_dst = lib.up(src), lib.up(np.zeros_like(src))
    lib.call("cyb_copy_blocks", lib.p(d_src), lib.p(d_dst), lib.p(lib.up(J)), hp(J), len(J), int(dtype == np.complex128), 0,
             lib.stream())
    out = d_dst.cpu().numpy()
    for b, o in zip(blocks, offs):
        assert np.array_equal(out[o:o + b.size].reshape([b.shape[p] for p in perm]), b.transpose(perm))   # bit-exact
    # conjugating copy
    if dtype == np.complex128:
        d_dst2 = lib.up(np.zeros_like(src))
        lib.call("cyb_copy_blocks", lib.p(d_src), lib.p(d_dst2), lib.p(lib.up(J)), hp(J), len(J), 1, 1, lib.stream())
        out = d_dst2.cpu().numpy()
        for b, o in zip(blocks, offs):
            assert np.array_equal(out[o:o + b.size].reshape([b.shape[p] for p in perm]), b.transpose(perm).conj())


def test_copy_blocks_box_into_slice(lib):
    """combine_legs-style: a permuted box written into a slice of a larger block, and back out."""
    rng = np.random.default_rng(3)
    small = rng.standard_normal((37, 5, 41))
    big = np.zeros((50, 300))
    # dst[b0:b0+37, c0:c0+205] = small.transpose(0,2,1).reshape(37, 205)   (F-style group)
    b0, c0 = 9, 60
    J = _jobs([0], [b0 * 300 + c0], np.asarray([[37, 5, 41]]), np.asarray([_cstr((37, 5, 41))]), np.asarray([[300, 1, 5]]))
    d_small, d_big = lib.up(small.reshape(-1)), lib.up(big.reshape(-1))
    lib.call("cyb_copy_blocks", lib.p(d_small), lib.p(d_big), lib.p(lib.up(J)), hp(J), 1, 0, 0, lib.stream())
    want = big.copy()
    want[b0:b0 + 37, c0:c0 + 205] = small.transpose(0, 2, 1).reshape(37, 205)
    got = d_big.cpu().numpy().reshape(50, 300)
    assert np.array_equal(got, want)
    # split_legs-style: slice of the big block -> contiguous small block
    J2 = _jobs([b0 * 300 + c0], [0], np.asarray([[37, 5, 41]]), np.asarray([[300, 1, 5]]), np.asarray([_cstr((37, 5, 41))]))
    d_back = lib.up(np.zeros(small.size))
    lib.call("cyb_copy_blocks", lib.p(d_big), lib.p(d_back), lib.p(lib.up(J2)), hp(J2), 1, 0, 0, lib.stream())
    assert np.array_equal(d_back.cpu().numpy().reshape(small.shape), small)


def _gemm_case(rng, dtype, problems):
    """problems: list of (M, N, [K...]).  Returns arenas and tables."""
    cx = dtype == np.complex128
    a_parts, b_parts, pairs, mn, pp, coff = [], [], [], [], [0], []
    ao = bo = co = 0

    def al(x):
        return (x + 15) // 16 * 16
    for (M, N, Ks) in problems:
        for K in Ks:
            A = rng.standard_normal((M, K)) + (1j * rng.standard_normal((M, K)) if cx else 0)
            B = rng.standard_normal((K, N)) + (1j * rng.standard_normal((K, N)) if cx else 0)
            a_parts.append((ao, A.astype(dtype)))
            b_parts.append((bo, B.astype(dtype)))
            pairs.append((ao, bo, K))
            ao += al(M * K)
            bo += al(K * N)
        mn.append((M, N))
        coff.append(co)
        co += al(M * N)
        pp.append(len(pairs))
    Aar, Bar = np.zeros(max(ao, 16), dtype=dtype), np.zeros(max(bo, 16), dtype=dtype)
    for o, m in a_parts:
        Aar[o:o + m.size] = m.reshape(-1)
    for o, m in b_parts:
        Bar[o:o + m.size] = m.reshape(-1)
    return Aar, Bar, max(co, 16), np.asarray(pairs, np.int64).reshape(-1, 3), np.asarray(mn, np.int64), np.asarray(pp, np.int64), np.asarray(coff, np.int64), a_parts, b_parts


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_grouped_gemm(lib, dtype):
    rng = np.random.default_rng(11)
    problems = [(1, 1, [1]), (5, 7, [3]), (17, 33, [9, 4, 1]), (128, 128, [64]), (130, 129, [31, 17]), (64, 300, [50]),
                (257, 66, [128, 3]), (3, 500, [7, 7, 7, 7]), (200, 1, [33]), (96, 96, [5, 6, 7, 8, 9]),
                # M <= 8: the streaming (HBM-bound) path, even / odd N, chunk edges, several pairs
                (8, 1025, [2, 3]), (4, 2048, [6]), (2, 513, [1]), (7, 512, [5, 1]), (1, 1300, [2]), (9, 1024, [3])]
    Aar, Bar, csize, pairs, mn, pp, coff, a_parts, b_parts = _gemm_case(rng, dtype, problems)
    d_c = lib.up(np.full(csize, np.nan, dtype=dtype))
    lib.call("cyb_grouped_gemm", lib.p(lib.up(Aar)), lib.p(lib.up(Bar)), lib.p(d_c), int(dtype == np.complex128),
             lib.p(lib.up(mn)), lib.p(lib.up(coff)), lib.p(lib.up(pp)), lib.p(lib.up(pairs)), lib.p(None), len(mn), lib.stream())
    out = d_c.cpu().numpy()
    q = 0
    for t, (M, N, Ks) in enumerate(problems):
        want = np.zeros((M, N), dtype=dtype)
        bound = np.zeros((M, N))
        for K in Ks:
            want += a_parts[q][1] @ b_parts[q][1]
            bound += np.abs(a_parts[q][1]) @ np.abs(b_parts[q][1])
            q += 1
        got = out[coff[t]:coff[t] + M * N].reshape(M, N)
        # float64 dot products: |err| <= (sum K) * eps * sum|a||b|; stated tolerance 1e-13 relative to that bound
        assert np.all(np.abs(got - want) <= 1e-13 * bound + 1e-300), (t, M, N, Ks)
    # subset launch: only blocks 3, 5 and 7 are written
    d_c2 = lib.up(np.full(csize, 7.0, dtype=dtype))
    sub = np.asarray([5, 3, 7], dtype=np.int64)
    lib.call("cyb_grouped_gemm", lib.p(lib.up(Aar)), lib.p(lib.up(Bar)), lib.p(d_c2), int(dtype == np.complex128),
             lib.p(lib.up(mn)), lib.p(lib.up(coff)), lib.p(lib.up(pp)), lib.p(lib.up(pairs)), lib.p(lib.up(sub)), 3, lib.stream())
    out2 = d_c2.cpu().numpy()
    for t, (M, N, Ks) in enumerate(problems):
        got = out2[coff[t]:coff[t] + M * N]
        if t in (3, 5, 7):
            assert np.allclose(got, out[coff[t]:coff[t] + M * N], rtol=0, atol=0)
        else:
            assert np.all(got == 7.0)


def _mats(rng, dtype, dims, kind="random"):
    cx = dtype == np.complex128
    mats = []
    for (m, n) in dims:
        A = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cx else 0)
        if kind == "lowrank" and min(m, n) > 2:
            r = max(1, min(m, n) // 2)
            A = (rng.standard_normal((m, r)) + (1j * rng.standard_normal((m, r)) if cx else 0)) @ \
                (rng.standard_normal((r, n)) + (1j * rng.standard_normal((r, n)) if cx else 0))
        elif kind == "zero":
            A = A * 0
        mats.append(np.ascontiguousarray(A.astype(dtype)))
    return mats


def _pack(mats, dtype):
    offs, o = [], 0
    for m in mats:
        offs.append(o)
        o += (m.size + 15) // 16 * 16
    ar = np.zeros(max(o, 16), dtype=dtype)
    for m, oo in zip(mats, offs):
        ar[oo:oo + m.size] = m.reshape(-1)
    return ar, np.asarray(offs, np.int64)


SVD_DIMS = [(1, 1), (2, 3), (5, 5), (17, 9), (9, 17), (40, 40), (64, 33), (100, 130), (150, 97), (1, 20), (20, 1),
            # >= 96 working rows: the block-Jacobi path (odd / even leading dimensions, ragged last block)
            (300, 300), (200, 331), (257, 131), (129, 640)]


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
@pytest.mark.parametrize("kind", ["random", "lowrank", "zero"])
def test_batched_svd(lib, dtype, kind):
    rng = np.random.default_rng(21)
    mats = _mats(rng, dtype, SVD_DIMS, kind)
    a, aoff = _pack(mats, dtype)
    ks = [min(m.shape) for m in mats]
    u, uoff = _pack([np.zeros((m.shape[0], k)) for m, k in zip(mats, ks)], dtype)
    s, soff = _pack([np.zeros(k) for k in ks], np.float64)
    vh, voff = _pack([np.zeros((k, m.shape[1])) for m, k in zip(mats, ks)], dtype)
    desc = np.asarray([(m.shape[0], m.shape[1], ao, uo, so, vo) for m, ao, uo, so, vo in zip(mats, aoff, uoff, soff, voff)], np.int64)
    d_u, d_s, d_vh, d_info = lib.up(u), lib.up(s), lib.up(vh), lib.empty(len(mats), torch.int64)
    lib.call("cyb_batched_svd", lib.p(lib.up(a)), lib.p(d_u), lib.p(d_s), lib.p(d_vh), desc.ctypes.data_as(C.c_void_p),
             len(mats), int(dtype == np.complex128), lib.p(d_info), lib.stream())
    U, S, Vh, info = d_u.cpu().numpy(), d_s.cpu().numpy(), d_vh.cpu().numpy(), d_info.cpu().numpy()
    assert np.all(info > 0), info
    for A, k, uo, so, vo in zip(mats, ks, uoff, soff, voff):
        m, n = A.shape
        Ui, Si, Vi = U[uo:uo + m * k].reshape(m, k), S[so:so + k], Vh[vo:vo + k * n].reshape(k, n)
        ref = np.linalg.svd(A, compute_uv=False)
        smax = max(ref.max(), 1e-300)
        assert np.all(np.abs(Si - ref) <= 1e-12 * smax)                          # 1e-12 of the largest singular value
        assert np.allclose(Ui @ np.diag(Si) @ Vi, A, rtol=0, atol=1e-12 * max(smax, 1) * max(m, n))
        assert np.allclose(Ui.conj().T @ Ui, np.eye(k), rtol=0, atol=1e-12 * max(m, n))
        assert np.allclose(Vi @ Vi.conj().T, np.eye(k), rtol=0, atol=1e-12 * max(m, n))


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
@pytest.mark.parametrize("kind", ["random", "lowrank", "zero"])
def test_batched_qr(lib, dtype, kind):
    rng = np.random.default_rng(22)
    mats = _mats(rng, dtype, SVD_DIMS, kind)
    a, aoff = _pack(mats, dtype)
    ks = [min(m.shape) for m in mats]
    q, qoff = _pack([np.zeros((m.shape[0], k)) for m, k in zip(mats, ks)], dtype)
    r, roff = _pack([np.zeros((k, m.shape[1])) for m, k in zip(mats, ks)], dtype)
    desc = np.asarray([(m.shape[0], m.shape[1], ao, qo, ro, 0) for m, ao, qo, ro in zip(mats, aoff, qoff, roff)], np.int64)
    d_q, d_r = lib.up(q), lib.up(r)
    lib.call("cyb_batched_qr", lib.p(lib.up(a)), lib.p(d_q), lib.p(d_r), desc.ctypes.data_as(C.c_void_p), len(mats),
             int(dtype == np.complex128), lib.stream())
    Q, R = d_q.cpu().numpy(), d_r.cpu().numpy()
    for A, k, qo, ro in zip(mats, ks, qoff, roff):
        m, n = A.shape
        Qi, Ri = Q[qo:qo + m * k].reshape(m, k), R[ro:ro + k * n].reshape(k, n)
        scale = max(np.abs(A).max(), 1.0)
        assert np.allclose(Qi @ Ri, A, rtol=0, atol=1e-12 * scale * max(m, n))
        assert np.allclose(Qi.conj().T @ Qi, np.eye(k), rtol=0, atol=1e-12 * max(m, n))
        assert np.all(np.tril(Ri, -1) == 0)
        # |diag R| is gauge invariant for full-rank input
        if kind == "random":
            ref = np.abs(np.diag(np.linalg.qr(A, mode="reduced")[1]))
            assert np.allclose(np.abs(np.diag(Ri)), ref, rtol=1e-10, atol=1e-12 * scale)


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_batched_eigh(lib, dtype):
    rng = np.random.default_rng(23)
    ns = [1, 2, 7, 33, 64, 100, 129, 300, 193]
    mats = []
    for n in ns:
        A = _mats(rng, dtype, [(n, n)])[0]
        Hm = A + A.conj().T
        if n == 33:
            Hm = Hm * 0 + np.diag(np.repeat([1.0, -1.0, 2.0], 11)).astype(dtype)   # degenerate +-lambda
        if n in (193, 300):
            # block-Jacobi path (the eigenvectors are read off the working copy, not off an accumulated factor):
            # n = 193: three eigenvalues of multiplicity 64 / 64 / 65 in a random basis; n = 300: a graded spectrum
            # 1 ... 1e-14 with random signs (the shift by 1.5 ||H||_F must keep every eigenvalue to 1e-12 ||H||)
            Qm, _ = np.linalg.qr(A)
            ev = (np.repeat([1.0, -1.0, 2.0], 65)[:n] if n == 193
                  else np.logspace(0, -14, n) * rng.choice([-1.0, 1.0], size=n))
            Hm = (Qm * ev) @ Qm.conj().T
            Hm = 0.5 * (Hm + Hm.conj().T)
        mats.append(np.ascontiguousarray(Hm))
    a, aoff = _pack(mats, dtype)
    w, woff = _pack([np.zeros(n) for n in ns], np.float64)
    v, voff = _pack([np.zeros((n, n)) for n in ns], dtype)
    desc = np.asarray([(n, n, ao, 0, wo, vo) for n, ao, wo, vo in zip(ns, aoff, woff, voff)], np.int64)
    d_w, d_v, d_info = lib.up(w), lib.up(v), lib.empty(len(ns), torch.int64)
    lib.call("cyb_batched_eigh", lib.p(lib.up(a)), lib.p(d_w), lib.p(d_v), desc.ctypes.data_as(C.c_void_p), len(ns),
             int(dtype == np.complex128), lib.p(d_info), lib.stream())
    W, V = d_w.cpu().numpy(), d_v.cpu().numpy()
    assert np.all(d_info.cpu().numpy() > 0)
    for Hm, n, wo, vo in zip(mats, ns, woff, voff):
        Wi, Vi = W[wo:wo + n], V[vo:vo + n * n].reshape(n, n)
        ref = np.linalg.eigvalsh(Hm)
        scale = max(np.abs(ref).max(), 1e-300)
        assert np.all(np.abs(Wi - ref) <= 1e-12 * scale * max(1, np.sqrt(n)))      # eigenvalues, ascending
        assert np.allclose(Vi @ np.diag(Wi) @ Vi.conj().T, Hm, rtol=0, atol=1e-12 * scale * n)
        assert np.allclose(Vi.conj().T @ Vi, np.eye(n), rtol=0, atol=1e-12 * n)


def test_fill_eye(lib):
    dims = np.asarray([(3, 5), (4, 2), (1, 1), (6, 6)], np.int64)
    offs = np.asarray([0, 16, 32, 48], np.int64)
    for dt, code in ((np.float64, 0), (np.complex128, 1)):
        d = lib.up(np.full(96, 9.0, dtype=dt))
        lib.call("cyb_fill_eye", lib.p(d), lib.p(lib.up(dims)), lib.p(lib.up(offs)), len(dims), code, lib.stream())
        out = d.cpu().numpy()
        for (m, n), o in zip(dims, offs):
            assert np.array_equal(out[o:o + m * n].reshape(m, n), np.eye(m, n, dtype=dt))


def test_truncate_select_golden(lib):
    """Known-answer vectors from the reference's _truncate_singular_values_selection."""
    z = np.load(__import__("pathlib").Path(__file__).parent / "golden" / "truncation.npz")
    for t in range(int(z["count"])):
        S, mask, args = z[f"t{t}/S"], z[f"t{t}/mask"], z[f"t{t}/args"]
        n = len(S)
        d_keep, d_res = lib.empty(n, torch.uint8), lib.empty(4, torch.float64)
        lib.call("cyb_truncate_select", lib.p(lib.up(S)), n, int(args[0]), int(args[1]), float(args[2]), float(args[3]),
                 float(args[4]), int(args[5]), lib.p(d_keep), lib.p(d_res), lib.stream())
        keep, res = d_keep.cpu().numpy()[:n].astype(bool), d_res.cpu().numpy()
        # identical kept set except at exactly degenerate cutoffs: compare the kept VALUES
        assert keep.sum() == mask.sum(), t
        assert np.array_equal(np.sort(S[keep]), np.sort(S[mask])), t
        assert abs(res[0] - float(z[f"t{t}/err"])) <= 1e-14, t
        assert abs(res[1] - float(z[f"t{t}/new_norm"])) <= 1e-14, t


@pytest.mark.parametrize("bad", [np.nan, np.inf])
def test_truncate_select_rejects_non_finite_input(lib, bad):
    """A NaN (unconverged / overflowed sector upstream) must not corrupt the rank sort: the kernel keeps nothing and
    flags the input (result[2] = -1); the host layer raises (next test)."""
    rng = np.random.default_rng(3)
    for n in (5, 300, 5000):
        S = np.abs(rng.standard_normal(n))
        S[rng.integers(0, n, size=max(1, n // 50))] = bad
        d_keep, d_res = lib.up(np.ones(n, np.uint8)), lib.empty(4, torch.float64)
        lib.call("cyb_truncate_select", lib.p(lib.up(S)), n, n // 2, 1, 0.0, 0.0, 1e-12, 1, lib.p(d_keep), lib.p(d_res), lib.stream())
        assert d_res.cpu().numpy()[2] == -1.0
        assert not d_keep.cpu().numpy()[:n].any()


def test_truncate_singular_values_raises_on_nan(backend):
    from cyten_b200._lib import CytenB200Error
    from cyten_b200.data import AbelianBackendData
    from cyten_b200.spaces import ElementarySpace, u1_symmetry
    leg = ElementarySpace(u1_symmetry, [[-1], [0], [2]], [3, 4, 2], False)
    bi = np.asarray([[0, 0], [1, 1], [2, 2]])
    blocks = [np.asarray([3.0, 2.0, 1.0]), np.asarray([5.0, np.nan, 1.0, 0.5]), np.asarray([2.5, 0.1])]
    S = AbelianBackendData.from_numpy_blocks(blocks, bi[:, :1], backend.device, dtype="float64", is_sorted=True)
    S.block_inds = bi
    with pytest.raises(CytenB200Error, match="non-finite"):
        backend.truncate_singular_values(S, leg, chi_max=4, svd_min=1e-12)


def test_blas1_helpers(lib):
    rng = np.random.default_rng(9)
    for dt, code in ((np.float64, 0), (np.complex128, 1)):
        x = (rng.standard_normal(1000) + (1j * rng.standard_normal(1000) if code else 0)).astype(dt)
        out = lib.empty(1, torch.float64)
        lib.call("cyb_sumsq", lib.p(lib.up(x)), len(x), code, lib.p(out), lib.stream())
        assert abs(float(out.cpu()[0]) - np.sum(np.abs(x) ** 2)) <= 1e-12 * np.sum(np.abs(x) ** 2)
        y = lib.up(np.ones_like(x))
        lib.call("cyb_axpby", lib.p(lib.up(x)), lib.p(y), len(x), code, 2.0, -1.0, lib.stream())
        assert np.allclose(y.cpu().numpy(), 2 * x - 1, rtol=1e-15, atol=1e-15)
    # gather rows / cols
    M = rng.standard_normal((6, 9))
    idx = np.asarray([0, 2, 5, 1, 4], np.int64)
    desc = np.asarray([(6, 9, 0, 0, 0, 3), (6, 9, 0, 32, 3, 2)], np.int64)
    d_dst = lib.up(np.zeros(96))
    lib.call("cyb_gather_rows_cols", lib.p(lib.up(M.reshape(-1))), lib.p(d_dst), lib.p(lib.up(desc)), lib.p(lib.up(idx)), 2, 0, 0, lib.stream())
    out = d_dst.cpu().numpy()
    assert np.array_equal(out[:27].reshape(3, 9), M[[0, 2, 5]])
    assert np.array_equal(out[32:50].reshape(2, 9), M[[1, 4]])
    d_dst = lib.up(np.zeros(96))
    lib.call("cyb_gather_rows_cols", lib.p(lib.up(M.reshape(-1))), lib.p(d_dst), lib.p(lib.up(desc)), lib.p(lib.up(idx)), 2, 1, 0, lib.stream())
    out = d_dst.cpu().numpy()
    assert np.array_equal(out[:18].reshape(6, 3), M[:, [0, 2, 5]])
    assert np.array_equal(out[32:44].reshape(6, 2), M[:, [1, 4]])
    # scale rows / cols
    dims, off, doff = np.asarray([(6, 9)], np.int64), np.asarray([0], np.int64), np.asarray([0], np.int64)
    d = rng.standard_normal(9)
    x = lib.up(M.reshape(-1).copy())
    lib.call("cyb_scale_rows_cols", lib.p(x), lib.p(lib.up(dims)), lib.p(lib.up(off)), lib.p(lib.up(d)), lib.p(lib.up(doff)), 1, 1, 0, lib.stream())
    assert np.array_equal(x.cpu().numpy().reshape(6, 9), M * d[None, :])


# ---- (f1) Krylov vector operations over whole arenas ------------------------------------------------------------
def _b1_arena(rng, sizes, dtype):
    """Blocks at 16-element aligned offsets with NaN padding (the kernels must never read the padding)."""
    offs, tot = [], 0
    for s in sizes:
        offs.append(tot)
        tot += (s + 15) // 16 * 16
    a = np.full(max(tot, 16), np.nan, dtype=dtype)
    for o, s in zip(offs, sizes):
        a[o:o + s] = rng.standard_normal(s) + (1j * rng.standard_normal(s) if dtype == np.complex128 else 0)
    return a, np.asarray(offs, np.int64)


def _b1_chunk_elems(lib, cx):
    if lib.kind == "cuda":
        from cyten_b200 import _lib
        return int(_lib.load().cyb_blas1_chunk_elems(int(cx)))
    return int(O.cyb_blas1_chunk_elems(int(cx)))


def _b1_chunks(sizes, chunk):
    c0 = np.zeros(len(sizes) + 1, np.int64)
    c0[1:] = np.cumsum((np.asarray(sizes, np.int64) + chunk - 1) // chunk)
    return c0


B1_SIZES = [1, 2, 3, 15, 16, 17, 511, 512, 513, 1023, 1024, 1025, 5000, 70001, 7]


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
@pytest.mark.parametrize("flags", [0, 1, 3])
def test_block_dot(lib, dtype, flags):
    rng = np.random.default_rng(31)
    x, ox = _b1_arena(rng, B1_SIZES, dtype)
    y, oy = _b1_arena(rng, B1_SIZES[::-1], dtype)
    oy = oy[::-1].copy()                                   # block t of x pairs with block t of y (other arena order)
    keep = [t for t in range(len(B1_SIZES)) if t != 4]     # one block without partner
    seg = np.stack([ox[keep], oy[keep], np.asarray(B1_SIZES, np.int64)[keep]], axis=1)
    c0 = _b1_chunks(seg[:, 2], _b1_chunk_elems(lib, dtype == np.complex128))
    out = lib.empty(2, torch.float64)
    work = lib.empty(148 * 8 * 2 + 8, torch.float64)
    for _ in range(2):                                     # the second call checks that the workspace re-arms itself
        lib.call("cyb_block_dot", lib.p(lib.up(x)), lib.p(lib.up(y)), lib.p(lib.up(seg)), lib.p(lib.up(c0)), len(seg),
                 int(c0[-1]), int(dtype == np.complex128), flags, lib.p(out), lib.p(work), lib.stream())
    ref = sum(np.sum((np.conj(x[a:a + n]) if flags & 1 else x[a:a + n]) * y[b:b + n]) for a, b, n in seg)
    got = out.cpu().numpy()
    scale = sum(np.sum(np.abs(x[a:a + n]) * np.abs(y[b:b + n])) for a, b, n in seg)
    assert abs(got[0] - np.real(ref)) <= 1e-14 * scale
    assert abs(got[1] - (0.0 if flags & 2 else np.imag(ref))) <= 1e-14 * scale


def test_block_dot_empty(lib):
    out = lib.up(np.ones(2))
    work = lib.empty(148 * 8 * 2 + 8, torch.float64)
    z = lib.up(np.zeros(16))
    lib.call("cyb_block_dot", lib.p(z), lib.p(z), lib.p(lib.up(np.zeros((1, 3), np.int64))), lib.p(lib.up(np.zeros(1, np.int64))),
             0, 0, 0, 1, lib.p(out), lib.p(work), lib.stream())
    assert np.array_equal(out.cpu().numpy(), [0.0, 0.0])


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
@pytest.mark.parametrize("with_dev,with_norm", [(False, False), (True, True)])
def test_block_lincomb(lib, dtype, with_dev, with_norm):
    rng = np.random.default_rng(32)
    cx = dtype == np.complex128
    x, ox = _b1_arena(rng, B1_SIZES, dtype)
    y, oy = _b1_arena(rng, B1_SIZES, dtype)
    z, oz = _b1_arena(rng, B1_SIZES, dtype)
    o, oo = _b1_arena(rng, B1_SIZES, dtype)
    sizes = np.asarray(B1_SIZES, np.int64)
    oy2, oz2 = oy.copy(), oz.copy()
    oy2[2] = -1                                            # y lacks block 2, z lacks blocks 5 and 6
    oz2[5:7] = -1
    seg = np.stack([ox, oy2, oz2, oo, sizes], axis=1)
    c0 = _b1_chunks(sizes, _b1_chunk_elems(lib, cx))
    coef = np.asarray([0.7, -0.2 if cx else 0.0, -1.3, 0.4 if cx else 0.0, 2.0, 0.0])
    dev = lib.up(np.asarray([0.35, 0.9 if cx else 0.0]))
    ptrs = (C.c_void_p * 3)(None, dev.data_ptr() if with_dev else None, None)
    d_o = lib.up(o)
    ss = lib.empty(1, torch.float64)
    work = lib.empty(148 * 8 * 2 + 8, torch.float64)
    lib.call("cyb_block_lincomb", lib.p(lib.up(x)), lib.p(lib.up(y)), lib.p(lib.up(z)), lib.p(d_o), lib.p(lib.up(seg)),
             lib.p(lib.up(c0)), len(seg), int(c0[-1]), int(cx), coef.ctypes.data_as(C.c_void_p), ptrs,
             lib.p(ss) if with_norm else None, lib.p(work), lib.stream())
    k = [complex(coef[0], coef[1]), complex(coef[2], coef[3]), complex(coef[4], coef[5])]
    if with_dev:
        k[1] *= complex(0.35, 0.9) if cx else 0.35
    if not cx:
        k = [c.real for c in k]
    got = d_o.cpu().numpy()
    tot = 0.0
    for t, n in enumerate(sizes):
        ref = k[0] * x[ox[t]:ox[t] + n]
        if oy2[t] >= 0:
            ref = ref + k[1] * y[oy[t]:oy[t] + n]
        if oz2[t] >= 0:
            ref = ref + k[2] * z[oz[t]:oz[t] + n]
        assert np.allclose(got[oo[t]:oo[t] + n], ref, rtol=0, atol=1e-14 * max(1.0, np.abs(ref).max()))
        pad = got[oo[t] + n:oo[t] + (n + 15) // 16 * 16]
        assert np.all(np.isnan(pad))                        # the padding of the result arena is not written
        tot += np.sum(np.abs(ref) ** 2)
    if with_norm:
        assert abs(ss.cpu().numpy()[0] - tot) <= 1e-13 * tot


@pytest.mark.parametrize("dtype,ddtype", [(np.float64, np.float64), (np.complex128, np.float64), (np.complex128, np.complex128)])
def test_scale_axis(lib, dtype, ddtype):
    rng = np.random.default_rng(33)
    shapes = [(3, 5, 4), (1, 7, 1), (40, 33, 9), (2, 1, 1500), (1, 1, 1), (17, 64, 1)]
    sizes = [int(np.prod(s)) for s in shapes]
    x, ox = _b1_arena(rng, sizes, dtype)
    d, od = _b1_arena(rng, [s[1] for s in shapes], ddtype)
    oo = ox[::-1].copy()                                  # output blocks in another order
    order = np.argsort(-np.arange(len(shapes)))
    out_sizes = [sizes[t] for t in order]
    _, ooff = _b1_arena(rng, out_sizes, dtype)
    oo = np.empty(len(shapes), np.int64)
    oo[order] = ooff
    desc = np.asarray([(ox[t], oo[t], s[0], s[1], s[2], od[t]) for t, s in enumerate(shapes)], np.int64)
    w0 = np.zeros(len(shapes) + 1, np.int64)
    chunk = _b1_chunk_elems(lib, dtype == np.complex128)
    w0[1:] = np.cumsum((np.asarray(sizes) + chunk - 1) // chunk)
    d_out = lib.empty(sum((s + 15) // 16 * 16 for s in sizes), torch.complex128 if dtype == np.complex128 else torch.float64)
    lib.call("cyb_scale_axis", lib.p(lib.up(x)), lib.p(d_out), lib.p(lib.up(desc)), lib.p(lib.up(w0)), len(shapes), int(w0[-1]),
             lib.p(lib.up(d)), int(dtype == np.complex128), int(ddtype == np.complex128), lib.stream())
    got = d_out.cpu().numpy()
    for t, s in enumerate(shapes):
        ref = x[ox[t]:ox[t] + sizes[t]].reshape(s) * d[od[t]:od[t] + s[1]][None, :, None]
        assert np.allclose(got[oo[t]:oo[t] + sizes[t]].reshape(s), ref, rtol=0, atol=1e-15 * max(1.0, np.abs(ref).max()))


def test_cutoff_inverse(lib):
    """cyb_cutoff_inverse == 1 / np.where(|x| < cutoff, inf, x) (numpy.cpp:645-656), bit for bit."""
    rng = np.random.default_rng(5)
    x = rng.standard_normal(1000)
    x[::7] = 1e-16
    x[3] = 0.0
    y = lib.up(np.zeros_like(x))
    lib.call("cyb_cutoff_inverse", lib.p(lib.up(x)), lib.p(y), len(x), 1e-12, lib.stream())
    with np.errstate(divide="ignore"):
        want = 1.0 / np.where(np.abs(x) < 1e-12, np.inf, x)
    assert np.array_equal(y.cpu().numpy(), want)

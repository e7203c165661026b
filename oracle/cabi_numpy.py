"""CPU oracle: a numpy restatement of every entry point of ``include/cyten_b200.h``.

TEST INFRASTRUCTURE ONLY -- nothing in the product package ``cyten_b200`` imports this module.
It is used by ``tests/`` (as the checker of the CUDA kernels and, on machines without a GPU, as the
stand-in library that lets the host-side logic of ``cyten_b200`` be exercised against the golden
vectors of the real reference), by ``__graft_entry__.smoke()`` and by ``bench.py``'s
``cpu_baseline`` leg.

Every function restates the reference algorithm its CUDA counterpart replaces and cites it
(paths relative to tenpy/cyten).  The dense linear algebra of the reference lives in third-party
libraries that are not under /root/reference: numpy >= 2.0 (``np.dot``, ``np.linalg.eigh``) and
scipy.linalg (``svd`` gesdd, ``qr``), both backed by OpenBLAS/LAPACK (pins: environment.yml:15-17);
the restatement calls exactly those routines, like ``NumpyBlockBackend`` does
(src/block_backend/numpy.cpp:658-680, 1219-1297).

Parity pinning: checked against golden vectors generated from the reference itself
(``tests/golden/*.npz`` made by ``tests/golden/make_golden.py``) in ``tests/test_oracle_golden.py``.

Pointers are plain integers (host addresses here): the functions have the C signatures, so the same
host code drives either this module or ``libcyten_b200.so``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import scipy.linalg

MAX_RANK = 8
JOB_WORDS = 3 + 3 * MAX_RANK
_last_error = b""


def _addr(p) -> int:
    if p is None:
        return 0
    if isinstance(p, C.c_void_p):
        return p.value or 0
    if isinstance(p, C.Array):
        return C.addressof(p)
    return int(p)


def _arr(p, n, dtype):
    """numpy view of ``n`` items of ``dtype`` at address ``p``."""
    a = _addr(p)
    n = int(n)
    if n == 0:
        return np.zeros(0, dtype=dtype)
    if a == 0:
        raise ValueError("NULL pointer")
    nbytes = n * np.dtype(dtype).itemsize
    buf = (C.c_char * nbytes).from_address(a)
    return np.frombuffer(buf, dtype=dtype, count=n)


def _i64(p, n):
    return _arr(p, n, np.int64)


def _dt(dtype: int):
    return np.complex128 if int(dtype) == 1 else np.float64


# ---- library ---------------------------------------------------------------------------------------
def cyb_version():
    return 100


def cyb_last_error():
    return _last_error


# ---- (a1) BlockInds: src/backends/block_inds.cpp ----------------------------------------------------
def cyb_lexsort_rows(rows, nrows, ncols, perm_out, stream=None):
    """BlockInds::lexsort_indices (block_inds.cpp:290-299): stable, last column primary."""
    r = _i64(rows, nrows * ncols).reshape(nrows, ncols)
    out = _i64(perm_out, nrows)
    out[:] = np.lexsort(r.T) if ncols and nrows else np.arange(nrows)
    return 0


def cyb_take_rows(rows, ncols, idx, nrows_out, cols, ncols_out, out, stream=None):
    """BlockInds::take / take_columns."""
    ix = _i64(idx, nrows_out) if _addr(idx) else np.arange(nrows_out)
    cs = _i64(cols, ncols_out) if _addr(cols) else np.arange(ncols_out)
    nsrc = int(ix.max()) + 1 if len(ix) else 0
    r = _i64(rows, nsrc * ncols).reshape(nsrc, ncols)
    _i64(out, nrows_out * ncols_out).reshape(nrows_out, ncols_out)[:] = r[ix][:, cs]
    return 0


def cyb_row_differences(rows, nrows, ncols, starts_out, count_out, stream=None):
    """BlockInds::find_row_differences(include_len=true) (block_inds.cpp:308-336)."""
    r = _i64(rows, nrows * ncols).reshape(nrows, ncols)
    starts = [0] if nrows else []
    for i in range(1, nrows):
        if np.any(r[i - 1] != r[i]):
            starts.append(i)
    starts.append(nrows)
    _i64(starts_out, len(starts))[:] = starts
    _i64(count_out, 1)[0] = len(starts)
    return 0


def cyb_pack_rows(rows, nrows, ncols, strides, keys_out, stream=None):
    """BlockInds::pack (block_inds.cpp:556-571)."""
    r = _i64(rows, nrows * ncols).reshape(nrows, ncols)
    _i64(keys_out, nrows)[:] = r @ _i64(strides, ncols)
    return 0


def cyb_common_sorted(a, na, b, nb, pairs_out, count_out, stream=None):
    """iter_common_sorted_1d merge walk (block_inds.cpp:592-626 with both lists strict)."""
    x, y = _i64(a, na), _i64(b, nb)
    out = _i64(pairs_out, 2 * min(na, nb)).reshape(-1, 2)
    i = j = n = 0
    while i < na and j < nb:
        if x[i] < y[j]:
            i += 1
        elif x[i] > y[j]:
            j += 1
        else:
            out[n] = (i, j)
            n += 1
            i += 1
            j += 1
    _i64(count_out, 1)[0] = n
    return 0


def cyb_fuse_charges(bi, nrows, ncols, cols, nlegs, sectors, sector_ptr, modulus, nsym, charge_out,
                     stream=None):
    """Symmetry::multiple_fusion_broadcast for U(1)/Z_N products (symmetry.cpp:375-402)."""
    if nrows == 0:
        return 0
    b = _i64(bi, nrows * ncols).reshape(nrows, ncols)
    cs, sp, mod = _i64(cols, nlegs), _i64(sector_ptr, nlegs), _i64(modulus, nsym)
    out = _i64(charge_out, nrows * nsym).reshape(nrows, nsym)
    out[:] = 0
    for f in range(nlegs):
        idx = b[:, cs[f]]
        nsec = int(idx.max()) + 1
        sec = _i64(_addr(sectors) + 8 * int(sp[f]) * nsym, nsec * nsym).reshape(nsec, nsym)
        out += sec[idx]
    for s in range(nsym):
        if mod[s] > 0:
            out[:, s] %= mod[s]
    return 0


# ---- (a3-a5) strided block copies ---------------------------------------------------------------------
def cyb_copy_blocks_workspace(jobs_host, njobs):
    return 8 * JOB_WORDS * int(njobs) + 8


def cyb_copy_blocks_prepare(jobs, njobs, workspace, workspace_bytes, stream=None):
    """Plan half of cyb_copy_blocks: here the 'schedule' is the job list itself."""
    if njobs:
        _i64(workspace, njobs * JOB_WORDS)[:] = _i64(jobs, njobs * JOB_WORDS)
    return 0


def cyb_copy_blocks_run(src, dst, workspace, njobs, dtype, conj, stream=None):
    return cyb_copy_blocks(src, dst, workspace, None, njobs, dtype, conj)


def cyb_copy_blocks(src, dst, jobs, jobs_host, njobs, dtype, conj, stream=None):
    """permute_axes / combine_legs set_item / split_legs get_item as strided box copies
    (abelian.cpp:988-1006, 1196-1217, 3418-3433)."""
    dt = _dt(dtype)
    J = _i64(jobs, njobs * JOB_WORDS).reshape(njobs, JOB_WORDS)
    isz = np.dtype(dt).itemsize
    for row in J:
        r = int(row[2])
        shape = tuple(int(x) for x in row[3:3 + r])
        if int(np.prod(shape)) == 0:
            continue
        ss = row[3 + MAX_RANK:3 + MAX_RANK + r]
        ds = row[3 + 2 * MAX_RANK:3 + 2 * MAX_RANK + r]
        s_ext = int(sum((n - 1) * abs(int(s)) for n, s in zip(shape, ss))) + 1
        d_ext = int(sum((n - 1) * abs(int(s)) for n, s in zip(shape, ds))) + 1
        s_flat = _arr(_addr(src) + isz * int(row[0]), s_ext, dt)
        d_flat = _arr(_addr(dst) + isz * int(row[1]), d_ext, dt)
        sv = np.lib.stride_tricks.as_strided(s_flat, shape, [int(s) * isz for s in ss])
        dv = np.lib.stride_tricks.as_strided(d_flat, shape, [int(s) * isz for s in ds], writeable=True)
        dv[...] = np.conj(sv) if conj else sv
    return 0


# ---- (a6) compose ---------------------------------------------------------------------------------------
def cyb_compose_plan_workspace(na, nb):
    return 8


def cyb_compose_plan(a_bi, na, ra, a_keep, b_bi, nb, rb, ncontr, contr_nsec, a_charge, b_charge, nsym,
                     a_mk, b_kn, a_off, b_off, align_elems, cap_out, cap_pairs, header, out_bi,
                     out_rowcol, out_mn, out_off, out_pair_ptr, pairs, workspace, workspace_bytes,
                     stream=None):
    """Planning part of abelian_compose_worker (abelian.cpp:1238-1469), step by step:
    F-style key over the contracted columns (:1270-1283), stable lexsort of A by (keep, key)
    (:1286-1301), grouping by keep columns (:1303-1335), charge match + merge-join (:1420-1459)."""
    A = _i64(a_bi, na * ra).reshape(na, ra)
    B = _i64(b_bi, nb * rb).reshape(nb, rb)
    nsec = _i64(contr_nsec, ncontr)
    qa = _i64(a_charge, na * nsym).reshape(na, nsym)
    qb = _i64(b_charge, nb * nsym).reshape(nb, nsym)
    amk = _i64(a_mk, 2 * na).reshape(na, 2)
    bkn = _i64(b_kn, 2 * nb).reshape(nb, 2)
    aoff, boff = _i64(a_off, na), _i64(b_off, nb)
    strides = np.ones(ncontr, dtype=np.int64)
    for c in range(1, ncontr):
        strides[c] = strides[c - 1] * nsec[c - 1]
    a_keys = A[:, a_keep:] @ strides[::-1]
    b_keys = B[:, :ncontr] @ strides
    a_sort = np.lexsort(np.hstack([a_keys[:, None], A[:, :a_keep]]).T)
    a_keepbi = A[a_sort][:, :a_keep]
    a_skeys = a_keys[a_sort]
    b_keepbi = B[:, ncontr:]

    def diffs(x):
        n = len(x)
        d = [0] + [i for i in range(1, n) if np.any(x[i - 1] != x[i])] + [n]
        return d
    a_sl, b_sl = diffs(a_keepbi), diffs(b_keepbi)
    lookup = {}
    for ga in range(len(a_sl) - 1):
        lookup.setdefault(tuple(qa[a_sort[a_sl[ga]]]), []).append(ga)
    rows, mn, rc, pp, prs = [], [], [], [0], []
    for gb in range(len(b_sl) - 1):
        b0, b1 = b_sl[gb], b_sl[gb + 1]
        for ga in lookup.get(tuple(qb[b0]), []):
            a0, a1 = a_sl[ga], a_sl[ga + 1]
            i, j, common = a0, b0, []
            while i < a1 and j < b1:
                if a_skeys[i] < b_keys[j]:
                    i += 1
                elif a_skeys[i] > b_keys[j]:
                    j += 1
                else:
                    common.append((i, j))
                    i += 1
                    j += 1
            if not common:
                continue
            arow = a_sort[a0]
            rows.append(np.concatenate([A[arow, :a_keep], B[b0, ncontr:]]))
            rc.append((arow, b0))
            mn.append((amk[arow, 0], bkn[b0, 1]))
            for (i, j) in common:
                prs.append((aoff[a_sort[i]], boff[j], amk[a_sort[i], 1]))
            pp.append(len(prs))
    n_out, n_pairs = len(rows), len(prs)
    fits = n_out <= cap_out and n_pairs <= cap_pairs
    H = _i64(header, 8)
    H[:] = 0
    H[0], H[1], H[3], H[5], H[6] = n_out, n_pairs, int(fits), len(a_sl) - 1, len(b_sl) - 1
    if fits:
        rr = a_keep + rb - ncontr
        if n_out:
            _i64(out_bi, n_out * rr).reshape(n_out, rr)[:] = np.asarray(rows).reshape(n_out, rr)
            _i64(out_rowcol, 2 * n_out).reshape(n_out, 2)[:] = rc
            _i64(out_mn, 2 * n_out).reshape(n_out, 2)[:] = mn
            sizes = np.asarray([m * n for m, n in mn], dtype=np.int64)
            padded = (sizes + align_elems - 1) // align_elems * align_elems
            _i64(out_off, n_out)[:] = np.concatenate([[0], np.cumsum(padded)[:-1]])
            H[2] = int(padded.sum())
            _i64(pairs, 3 * n_pairs).reshape(n_pairs, 3)[:] = prs
            H[4] = int(sum(mn[t][0] * mn[t][1] * prs[q][2] for t in range(n_out) for q in range(pp[t], pp[t + 1])))
        _i64(out_pair_ptr, n_out + 1)[:] = pp
    return 0


def cyb_grouped_gemm(a, b, c, dtype, out_mn, out_off, out_pair_ptr, pairs, block_subset, nblocks,
                     stream=None):
    """matrix_dot + operator+ loop (abelian.cpp:1436-1446; np.dot, numpy.cpp:1219-1225)."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    sub = _i64(block_subset, nblocks) if _addr(block_subset) else np.arange(nblocks)
    nmax = int(sub.max()) + 1 if nblocks else 0
    mn = _i64(out_mn, 2 * nmax).reshape(nmax, 2)
    off = _i64(out_off, nmax)
    pp = _i64(out_pair_ptr, nmax + 1)
    pr = _i64(pairs, 3 * int(pp[nmax] if nmax else 0)).reshape(-1, 3)
    for t in sub:
        M, N = int(mn[t, 0]), int(mn[t, 1])
        acc = None
        for q in range(pp[t], pp[t + 1]):
            K = int(pr[q, 2])
            Am = _arr(_addr(a) + isz * int(pr[q, 0]), M * K, dt).reshape(M, K)
            Bm = _arr(_addr(b) + isz * int(pr[q, 1]), K * N, dt).reshape(K, N)
            acc = np.dot(Am, Bm) if acc is None else acc + np.dot(Am, Bm)
        if acc is None:
            acc = np.zeros((M, N), dtype=dt)
        _arr(_addr(c) + isz * int(off[t]), M * N, dt).reshape(M, N)[:] = acc
    return 0


def cyb_grouped_gemm_workspace(nblocks):
    return 8 * int(nblocks) + 8


def cyb_grouped_gemm_prepare(dtype, out_mn, out_pair_ptr, pairs, block_subset, nblocks, workspace,
                             workspace_bytes, stream=None):
    """Plan half of cyb_grouped_gemm: the schedule of the oracle is just the list of blocks to do."""
    sub = _i64(block_subset, nblocks) if _addr(block_subset) else np.arange(nblocks)
    if nblocks:
        _i64(workspace, nblocks)[:] = sub
    return 0


def cyb_grouped_gemm_skinny_rows():
    return 8


def cyb_grouped_gemm_run(a, b, c, dtype, out_mn, out_off, out_pair_ptr, pairs, nblocks, workspace, kernels=3,
                         stream=None):
    return cyb_grouped_gemm(a, b, c, dtype, out_mn, out_off, out_pair_ptr, pairs, workspace, nblocks)


# ---- (a7, a9, a10) factorisations --------------------------------------------------------------------------
def cyb_batched_svd(a, u, s, vh, desc_host, nmat, dtype, info, stream=None):
    """NumpyBlockBackend::matrix_svd (numpy.cpp:1247-1297): scipy.linalg.svd(full_matrices=False),
    gesdd with gesvd fallback."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    D = _i64(desc_host, 6 * nmat).reshape(nmat, 6)
    for i, (m, n, ao, uo, so, vo) in enumerate(D):
        m, n = int(m), int(n)
        k = min(m, n)
        if k == 0:
            continue
        A = _arr(_addr(a) + isz * int(ao), m * n, dt).reshape(m, n)
        try:
            U, S, Vh = scipy.linalg.svd(A, full_matrices=False, lapack_driver="gesdd")
        except np.linalg.LinAlgError:
            U, S, Vh = scipy.linalg.svd(A, full_matrices=False, lapack_driver="gesvd")
        _arr(_addr(u) + isz * int(uo), m * k, dt).reshape(m, k)[:] = U
        _arr(_addr(s) + 8 * int(so), k, np.float64)[:] = S
        _arr(_addr(vh) + isz * int(vo), k * n, dt).reshape(k, n)[:] = Vh
        if _addr(info):
            _i64(info, nmat)[i] = 1
    return 0


def cyb_batched_qr(a, q, r, desc_host, nmat, dtype, stream=None):
    """NumpyBlockBackend::matrix_qr(full=false) (numpy.cpp:1236-1245): scipy.linalg.qr economic."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    D = _i64(desc_host, 6 * nmat).reshape(nmat, 6)
    for (m, n, ao, qo, ro, _) in D:
        m, n = int(m), int(n)
        k = min(m, n)
        if k == 0:
            continue
        A = _arr(_addr(a) + isz * int(ao), m * n, dt).reshape(m, n)
        Q, R = scipy.linalg.qr(A, mode="economic")
        _arr(_addr(q) + isz * int(qo), m * k, dt).reshape(m, k)[:] = Q
        _arr(_addr(r) + isz * int(ro), k * n, dt).reshape(k, n)[:] = R
    return 0


def cyb_batched_eigh(a, w, v, desc_host, nmat, dtype, info, stream=None):
    """NumpyBlockBackend::eigh (numpy.cpp:658-680): np.linalg.eigh, ascending."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    D = _i64(desc_host, 6 * nmat).reshape(nmat, 6)
    for i, (m, n, ao, _, wo, vo) in enumerate(D):
        n = int(n)
        if n == 0:
            continue
        A = _arr(_addr(a) + isz * int(ao), n * n, dt).reshape(n, n)
        W, V = np.linalg.eigh(A)
        _arr(_addr(w) + 8 * int(wo), n, np.float64)[:] = W
        _arr(_addr(v) + isz * int(vo), n * n, dt).reshape(n, n)[:] = V
        if _addr(info):
            _i64(info, nmat)[i] = 1
    return 0


def cyb_fill_eye(dst, dims, off, nmat, dtype, stream=None):
    """eye_matrix + slice for absent sectors (abelian.cpp:3524-3538), eye_data (:1790-1826)."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    d = _i64(dims, 2 * nmat).reshape(nmat, 2)
    o = _i64(off, nmat)
    for (m, n), oo in zip(d, o):
        if m * n:
            _arr(_addr(dst) + isz * int(oo), int(m * n), dt).reshape(int(m), int(n))[:] = np.eye(int(m), int(n), dtype=dt)
    return 0


# ---- (a8) truncation -----------------------------------------------------------------------------------------
def truncate_selection(S, chi_max=None, chi_min=1, degeneracy_tol=0.0, trunc_cut=0.0, svd_min=None,
                       minimize_error=True):
    """TensorBackend::_truncate_singular_values_selection (tensor_backend.cpp:138-242), qdims=None."""
    S = np.asarray(S, dtype=np.float64)
    merr = S ** 2
    piv = np.argsort(merr, kind="stable")
    S, merr = S[piv], merr[piv]
    logS = np.log(np.choose(S <= 1.0e-100, [S, 1.0e-100 * np.ones(len(S))]))
    n = len(S)
    good = np.ones(n, dtype=bool)

    def combine(good, good2):
        res = good & good2
        return res if np.any(res) else good
    if chi_max is not None and chi_max < n:
        g2 = np.zeros(n, dtype=bool)
        g2[-chi_max:] = True
        good = combine(good, g2)
    if chi_min > 1:
        g2 = np.ones(n, dtype=bool)
        g2[-chi_min + 1:] = False
        good = combine(good, g2)
    if degeneracy_tol > 0:
        g2 = np.empty(n, dtype=bool)
        g2[0] = True
        g2[1:] = np.greater_equal(logS[1:] - logS[:-1], degeneracy_tol)
        good = combine(good, g2)
    if svd_min is not None:
        good = combine(good, np.greater_equal(S, svd_min))
    good = combine(good, np.cumsum(merr) > trunc_cut * trunc_cut)
    nz = np.nonzero(good)[0]
    cut = nz[0] if minimize_error else nz[-1]
    err, new_norm = np.sum(merr[:cut]), np.sum(merr[cut:])
    mask = np.zeros(n, dtype=bool)
    np.put(mask, piv[cut:], True)
    return mask, float(err), float(new_norm)


def cyb_truncate_select(s, n, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error,
                        keep, result, stream=None):
    S = _arr(s, n, np.float64)
    if not np.all(np.isfinite(S)):          # same contract as the CUDA kernel: no selection, result[2] = -1
        _arr(keep, n, np.uint8)[:] = 0
        _arr(result, 4, np.float64)[:] = (0.0, 0.0, -1.0, 0.0)
        return 0
    mask, err, new_norm = truncate_selection(S, None if chi_max < 0 else int(chi_max), int(chi_min),
                                             float(degeneracy_tol), float(trunc_cut),
                                             None if svd_min < 0 else float(svd_min), bool(minimize_error))
    _arr(keep, n, np.uint8)[:] = mask
    r = _arr(result, 4, np.float64)
    r[0], r[1], r[2], r[3] = err, new_norm, mask.sum(), err + new_norm
    return 0


def cyb_truncate_select_static(s, n, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error,
                               keep, result, stream=None):
    """The no-allocation variant has the same results; its size limit is CYB_TRUNCATE_STATIC_MAX_N of include/cyten_b200.h."""
    if n > 65536:
        return 3          # CYB_ERR_LIMIT
    return cyb_truncate_select(s, n, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error, keep, result, stream)


# ---- BLAS-1 helpers ---------------------------------------------------------------------------------------------
def cyb_sumsq(x, n, dtype, out, stream=None):
    v = _arr(x, n, _dt(dtype))
    _arr(out, 1, np.float64)[0] = float(np.sum(np.abs(v) ** 2))
    return 0


def cyb_axpby(x, y, n, dtype, alpha, beta, stream=None):
    xv = _arr(x, n, _dt(dtype)).copy()
    yv = _arr(y, n, _dt(dtype))
    yv[:] = alpha * xv if beta == 0.0 else alpha * xv + beta * yv
    return 0


def cyb_middle_apply_max_rows():
    return 16


def cyb_middle_apply_unit_cols():
    return 256


def cyb_middle_apply(inp, out, groups, desc, coef, work0, ngroups, nwork, max_rows, dtype, stream=None):
    """out[o][a][b] = sum_i C[o][i] in[i][a][b] per group (see include/cyten_b200.h)."""
    dt = _dt(dtype)
    G = _i64(groups, 6 * ngroups).reshape(ngroups, 6)
    if ngroups == 0:
        return 0
    ndesc = int(max(g[4] + g[2] + g[3] for g in G))
    D = _i64(desc, 2 * ndesc).reshape(ndesc, 2)
    ncoef = int(max(g[5] + g[2] * g[3] for g in G))
    Cf = _arr(coef, ncoef, dt)
    hi = 0
    for na, nb, nin, nout, d0, c0 in G:
        for off, st in D[d0:d0 + nin + nout]:
            hi = max(hi, int(off + (na - 1) * st + nb))
    X = _arr(inp, hi, dt)
    Y = _arr(out, hi, dt)
    for na, nb, nin, nout, d0, c0 in G:
        C = Cf[c0:c0 + nin * nout].reshape(nout, nin)
        rows = np.stack([np.stack([X[off + a * st: off + a * st + nb] for a in range(na)]) for off, st in D[d0:d0 + nin]]) \
            if nin else np.zeros((0, na, nb), dt)
        res = np.tensordot(C, rows, axes=(1, 0))
        for o, (off, st) in enumerate(D[d0 + nin:d0 + nin + nout]):
            for a in range(na):
                Y[off + a * st: off + a * st + nb] = res[o, a]
    return 0


def cyb_cutoff_inverse(x, y, n, cutoff, stream=None):
    """1 / np.where(np.abs(a) < cutoff, np.inf, a)  (NumpyBlockBackend::cutoff_inverse, numpy.cpp:645-656)."""
    xv = _arr(x, n, np.float64).copy()
    yv = _arr(y, n, np.float64)
    with np.errstate(divide="ignore"):
        yv[:] = 1.0 / np.where(np.abs(xv) < cutoff, np.inf, xv)
    return 0


def cyb_scale_rows_cols(x, dims, off, d, d_off, nmat, axis, dtype, stream=None):
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    D = _i64(dims, 2 * nmat).reshape(nmat, 2)
    O, DO = _i64(off, nmat), _i64(d_off, nmat)
    for (m, n), o, do in zip(D, O, DO):
        M = _arr(_addr(x) + isz * int(o), int(m * n), dt).reshape(int(m), int(n))
        dd = _arr(_addr(d) + 8 * int(do), int(m if axis == 0 else n), np.float64)
        M *= dd[:, None] if axis == 0 else dd[None, :]
    return 0


def cyb_gather_rows_cols(src, dst, desc, idx, nmat, axis, dtype, stream=None):
    """apply_mask along one axis (abelian.cpp:2483-2582)."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    D = _i64(desc, 6 * nmat).reshape(nmat, 6)
    for (m, n, so, do, ib, ic) in D:
        m, n, ic = int(m), int(n), int(ic)
        ix = _i64(_addr(idx) + 8 * int(ib), ic)
        S = _arr(_addr(src) + isz * int(so), m * n, dt).reshape(m, n)
        if axis == 0:
            _arr(_addr(dst) + isz * int(do), ic * n, dt).reshape(ic, n)[:] = S[ix, :]
        else:
            _arr(_addr(dst) + isz * int(do), m * ic, dt).reshape(m, ic)[:] = S[:, ix]
    return 0


# ---- (f1) Krylov vector operations over whole arenas --------------------------------------------------------
def cyb_blas1_workspace_bytes():
    return 64


def cyb_blas1_chunk_elems(dtype):
    return 512 if int(dtype) == 1 else 1024


def cyb_block_dot(x, y, seg, chunk0, nseg, nchunk, dtype, conj_x, out, work, stream=None):
    """inner (abelian.cpp:2159-2211): sum over the common blocks of block_backend.inner = sum conj(a) * b."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    o = _arr(out, 2, np.float64)
    tot = 0.0 + 0.0j
    for ox, oy, n in _i64(seg, 3 * nseg).reshape(nseg, 3):
        a = _arr(_addr(x) + isz * int(ox), int(n), dt)
        b = _arr(_addr(y) + isz * int(oy), int(n), dt)
        tot += np.sum((np.conj(a) if int(conj_x) & 1 else a) * b)
    o[0], o[1] = np.real(tot), (0.0 if int(conj_x) & 2 else np.imag(tot))
    return 0


def cyb_block_lincomb(x, y, z, out, seg, chunk0, nseg, nchunk, dtype, coef, coef_dev, sumsq, work, stream=None):
    """linear_combination / mul (abelian.cpp:2253-2301, 2759-2778): per block a*v + b*w, a*v where w has no such
    block, b*w where v has none; third operand for the Lanczos three-term recurrence."""
    dt = _dt(dtype)
    isz = np.dtype(dt).itemsize
    k = _arr(coef, 6, np.float64)
    k = [complex(k[0], k[1]), complex(k[2], k[3]), complex(k[4], k[5])]
    if _addr(coef_dev):
        ptrs = _arr(coef_dev, 3, np.uint64)
        for i in range(3):
            if int(ptrs[i]):
                d = _arr(int(ptrs[i]), 2, np.float64)
                k[i] *= complex(d[0], d[1]) if dt is np.complex128 else d[0]
    if dt is np.float64:
        k = [c.real for c in k]
    ss = 0.0
    for row in _i64(seg, 5 * nseg).reshape(nseg, 5):
        n = int(row[4])
        r = np.zeros(n, dtype=dt)
        for i, base in enumerate((x, y, z)):
            if _addr(base) and row[i] >= 0:
                r = r + k[i] * _arr(_addr(base) + isz * int(row[i]), n, dt)
        _arr(_addr(out) + isz * int(row[3]), n, dt)[:] = r
        ss += float(np.sum(np.abs(r) ** 2))
    if _addr(sumsq):
        _arr(sumsq, 1, np.float64)[0] = ss
    return 0


def cyb_scale_axis(inp, out, desc, work0, nblk, nwork, diag, dtype, diag_dtype, stream=None):
    """scale_axis (abelian.cpp:3177-3232): every block times the diagonal block of its sector along one axis."""
    dt, ddt = _dt(dtype), _dt(diag_dtype)
    isz, dsz = np.dtype(dt).itemsize, np.dtype(ddt).itemsize
    for oi, oo, outer, dleg, inner, doff in _i64(desc, 6 * nblk).reshape(nblk, 6):
        n = int(outer * dleg * inner)
        a = _arr(_addr(inp) + isz * int(oi), n, dt).reshape(int(outer), int(dleg), int(inner))
        d = _arr(_addr(diag) + dsz * int(doff), int(dleg), ddt)
        _arr(_addr(out) + isz * int(oo), n, dt).reshape(a.shape)[:] = a * d[None, :, None]
    return 0


def call(name: str, *args):
    """Same contract as ``cyten_b200._lib.call``."""
    rc = globals()[name](*args)
    if rc != 0:
        raise RuntimeError(f"oracle {name} failed with {rc}")

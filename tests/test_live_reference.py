"""Parity against the UNMODIFIED reference run live (oracle/_ref, built by oracle/build_ref.py; it travels to the GPU
box as a built artefact).  Random tensors are made by the reference's own ``SymmetricTensor.from_random_normal`` /
``randomly_drop_blocks``, pushed through the same operation on both sides, and compared in the reference's
tensor format: integer tables bit-exact, data movement bit-exact, contractions and factorisations to the
tolerance stated in each test.  Skipped when the reference has not been built.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
REF = ROOT / "oracle" / "_ref"
if not list((REF / "cyten").glob("_core*.so")):
    pytest.skip("oracle/_ref is not built (python oracle/build_ref.py)", allow_module_level=True)
sys.path.insert(0, str(REF))
import cyten as rc  # noqa: E402

from cyten_b200 import tensors as ct  # noqa: E402
from cyten_b200.interop import diagonal_to_numpy, from_reference, to_reference  # noqa: E402
from cyten_b200.spaces import AbelianSymmetry  # noqa: E402

SYMS = {"u1": (lambda: rc.u1_symmetry, (0,)),
        "u1u1": (lambda: rc.u1_symmetry * rc.u1_symmetry, (0, 0)),
        "z4": (lambda: rc.z4_symmetry, (4,))}
CASES = [("u1", "float64", 96), ("u1", "complex128", 64), ("u1u1", "float64", 80), ("z4", "complex128", 48)]


def _leg(sym, moduli, rng, dim, nsec, is_dual=False):
    secs = set()
    nsec = min(nsec, int(np.prod([m if m else 3 for m in moduli])))
    while len(secs) < nsec:
        secs.add(tuple(int(rng.integers(0, m)) if m else int(rng.integers(-1, 2)) for m in moduli))
    secs = np.asarray(sorted(secs), dtype=np.int64).reshape(len(secs), -1)
    secs = secs[np.lexsort(secs.T)]      # the reference's sector order (last column primary): trivial basis_perm
    mult = rng.multinomial(dim - len(secs), np.ones(len(secs)) / len(secs)) + 1
    return rc.ElementarySpace.from_defining_sectors(sym, secs, mult, is_dual=is_dual)


def _setup(name, dtype, chi, seed):
    mk, moduli = SYMS[name]
    sym = mk()
    rng = np.random.default_rng(seed)
    np.random.seed(seed)
    v = _leg(sym, moduli, rng, chi, 5)
    p = _leg(sym, moduli, rng, 3, 2)
    rdt = rc.Dtype.complex128 if dtype == "complex128" else rc.Dtype.float64
    be = rc.get_backend("abelian", "numpy")
    A = rc.SymmetricTensor.from_random_normal([v, p], [v], backend=be, labels=["vL", "p", "vR"], dtype=rdt)
    B = rc.SymmetricTensor.from_random_normal([v, p], [v], backend=be, labels=["vL", "p", "vR"], dtype=rdt)
    return sym, AbelianSymmetry(moduli), A, B


def _num(x):
    """Reference Scalar -> python number."""
    x = x.to_numpy() if hasattr(x, "to_numpy") else x
    return complex(x) if np.iscomplexobj(x) else float(x)


def _same(ours, ref, tol=0.0):
    """ours (arena tensor) == ref (reference tensor): tables bit-exact, blocks within tol * max|ref|."""
    bi = np.asarray(ref.data.block_inds.to_numpy()).reshape(len(ref.data.blocks), ours.data.block_inds.shape[1])
    assert np.array_equal(ours.data.block_inds, bi)
    rb = [np.asarray(b.to_numpy()) for b in ref.data.blocks]
    scale = max([np.abs(b).max() for b in rb if b.size] + [1e-300])
    for o, r in zip(ours.data.to_numpy_blocks(), rb):
        assert o.shape == r.shape
        if tol == 0.0:
            assert np.array_equal(o, r)
        else:
            assert np.abs(o - r).max() <= tol * scale


@pytest.mark.parametrize("name,dtype,chi", CASES)
def test_roundtrip_and_data_movement(backend, name, dtype, chi):
    sym, msym, A, B = _setup(name, dtype, chi, 1)
    a = from_reference(A, backend, msym)
    _same(a, A)
    back = to_reference(a, rc, sym)
    assert rc.almost_equal(back, A, rtol=0, atol=0)
    # permute_legs, combine_legs, split_legs: bit-exact data movement
    _same(ct.permute_legs(a, ["vR", "p"], ["vL"]), rc.permute_legs(A, ["vR", "p"], ["vL"], bend_right=None))
    ac, Ac = ct.combine_legs(a, [0, 1]), rc.combine_legs(A, [0, 1])
    _same(ac, Ac)
    _same(ct.split_legs(ac), rc.split_legs(Ac))


@pytest.mark.parametrize("name,dtype,chi", CASES)
def test_tdot_and_compose(backend, name, dtype, chi):
    sym, msym, A, B = _setup(name, dtype, chi, 2)
    a, b = from_reference(A, backend, msym), from_reference(B, backend, msym)
    _same(ct.tdot(a, b, [2], [0]), rc.tdot(A, B, ["vR"], ["vL"]), tol=1e-13)


@pytest.mark.parametrize("name,dtype,chi", CASES)
def test_svd_truncated_svd_qr(backend, name, dtype, chi):
    sym, msym, A, B = _setup(name, dtype, chi, 3)
    TH = rc.tdot(A, B, ["vR"], ["vL"])          # two-site theta [vL, p | p, vR], rank-deficient sectors included
    th = from_reference(TH, backend, msym)
    U, S, Vh = rc.svd(TH)
    u, s, vh = ct.svd(th)
    s_ref = diagonal_to_numpy(S)
    assert np.abs(diagonal_to_numpy(s) - s_ref).max() <= 1e-12 * s_ref.max()      # singular values
    assert np.array_equal(u.data.block_inds, np.asarray(U.data.block_inds.to_numpy()).reshape(len(U.data.blocks), -1))
    rec = ct.compose(ct.compose(u, _diag(s, backend)), vh)                          # U S Vh = theta (vectors: gauge)
    _same(rec, TH, tol=1e-12)
    chi_max = max(2, int(0.6 * chi))
    Ut, St, Vt, err, ren = rc.truncated_svd(TH, chi_max=chi_max, svd_min=1e-10)
    ut, st, vt, err2, ren2 = ct.truncated_svd(th, chi_max=chi_max, svd_min=1e-10)
    assert np.array_equal(np.asarray(St.leg.multiplicities), st.leg.multiplicities)   # kept sector dimensions
    assert np.array_equal(np.asarray(St.leg.defining_sectors).reshape(st.leg.defining_sectors.shape), st.leg.defining_sectors)
    assert abs(err - err2) <= 1e-12
    Q, R = rc.qr(A)
    q, r = ct.qr(from_reference(A, backend, msym))
    _same(ct.compose(q, r), A, tol=1e-12)


def _diag(S, be):
    from cyten_b200.tensors import SymmetricTensor
    blocks = [np.diag(b) for b in S.data.to_numpy_blocks()]
    return SymmetricTensor.from_numpy_blocks(blocks, S.data.block_inds, [S.leg], [S.leg], be, labels=["b", "c"],
                                             dtype=S.data.dtype)


# ---- (f1) Krylov vector operations and the Lanczos loop against the reference's own ------------------------------------
@pytest.mark.parametrize("name,dtype,chi", CASES)
def test_vector_ops(backend, name, dtype, chi):
    from cyten.testing import randomly_drop_blocks
    sym, msym, A, B = _setup(name, dtype, chi, 5)
    B2 = randomly_drop_blocks(1.0 * B, empty_ok=False, max_blocks=max(1, len(B.data.blocks) // 2),
                              np_random=np.random.default_rng(3))
    a, b, b2 = (from_reference(T, backend, msym) for T in (A, B, B2))
    scale = _num(rc.norm(A)) * _num(rc.norm(B))
    for x, X, y, Y in ((a, A, b, B), (a, A, b2, B2), (b2, B2, a, A)):
        assert abs(ct.inner(x, y) - _num(rc.inner(X, Y))) <= 1e-14 * scale                 # inner
    assert abs(ct.norm(a) - _num(rc.norm(A))) <= 1e-14 * _num(rc.norm(A))                     # norm
    Bd, bd = rc.dagger(B), ct.dagger(b)
    _same(bd, Bd)                                                                              # dagger: bit-exact
    assert bd.labels == list(Bd.labels)
    assert abs(ct.inner(a, bd, do_dagger=False) - _num(rc.inner(A, Bd, do_dagger=False))) <= 1e-14 * scale
    co = (0.3 - 0.2j, 1.7) if dtype == "complex128" else (0.3, -1.7)
    lin = rc.tensors.linear_combination
    _same(ct.linear_combination(co[0], a, co[1], b2), lin(co[0], A, co[1], B2), tol=1e-15)        # union of the blocks
    _same(ct.linear_combination(co[0], b2, co[1], a), lin(co[0], B2, co[1], A), tol=1e-15)
    _same(co[0] * a, co[0] * A, tol=1e-15)                                                     # mul
    _same(a - b, A - B, tol=1e-15)


@pytest.mark.parametrize("name,dtype,chi", CASES)
def test_scale_axis(backend, name, dtype, chi):
    from cyten_b200.interop import diagonal_from_reference
    sym, msym, A, B = _setup(name, dtype, chi, 6)
    U, S, Vh = rc.svd(rc.tdot(A, B, ["vR"], ["vL"]))
    sd = diagonal_from_reference(S, backend, msym)
    _same(ct.scale_axis(from_reference(U, backend, msym), sd, 2), rc.scale_axis(U, S, 2), tol=1e-15)
    _same(ct.scale_axis(from_reference(Vh, backend, msym), sd, 0), rc.scale_axis(Vh, S, 0), tol=1e-15)
    P = rc.permute_legs(U, [0, 2], [1], bend_right=None)          # the scaled leg in the middle of the block axes
    _same(ct.scale_axis(from_reference(P, backend, msym), sd, 1), rc.scale_axis(P, S, 1), tol=1e-15)


@pytest.mark.parametrize("name,dtype,chi", CASES)
@pytest.mark.parametrize("opts", [{}, {"N_cache": 3, "N_max": 30}, {"reortho": True, "N_max": 25}])
def test_lanczos_matches_reference_lanczos(backend, name, dtype, chi, opts):
    """The same Hermitian operator theta -> K theta + theta K2 and the same start vector through
    cyten.krylov_based.lanczos and cyten_b200.krylov.lanczos: same number of iterations, same Ritz value, same vector."""
    from cyten import krylov_based, sparse
    from cyten_b200.krylov import lanczos
    sym, msym, A, B = _setup(name, dtype, chi // 4, 7)
    TH = rc.tdot(A, B, ["vR"], ["vL"])
    rdt = rc.Dtype.complex128 if dtype == "complex128" else rc.Dtype.float64
    rb = rc.get_backend("abelian", "numpy")
    X = rc.SymmetricTensor.from_random_normal(TH.codomain, TH.codomain, backend=rb, dtype=rdt)
    X2 = rc.SymmetricTensor.from_random_normal(TH.domain, TH.domain, backend=rb, dtype=rdt)
    K, K2 = X + X.hc, X2 + X2.hc

    class RefOp(sparse.LinearOperator):
        def matvec(self, t):
            return rc.compose(K, t) + rc.compose(t, K2)

    class Op:
        def __init__(self):
            self.k, self.k2 = from_reference(K, backend, msym), from_reference(K2, backend, msym)

        def matvec(self, t):
            return ct.compose(self.k, t) + ct.compose(t, self.k2)

    Er, psir, Nr = krylov_based.lanczos(RefOp(), TH, dict(opts))
    E, psi, N = lanczos(Op(), from_reference(TH, backend, msym), dict(opts))
    assert N == Nr
    assert abs(E - _num(Er)) <= 1e-12 * (abs(_num(Er)) + 1.0)
    assert abs(abs(ct.inner(from_reference(psir, backend, msym), psi)) - 1.0) <= 1e-10


@pytest.mark.parametrize("name,dtype,chi", CASES[:2])
@pytest.mark.parametrize("delta", [-0.05j, -0.1, 0.03 - 0.02j])
def test_lanczos_evolution_matches_reference(backend, name, dtype, chi, delta):
    """exp(delta H) psi through cyten.krylov_based.LanczosEvolution and cyten_b200.krylov.LanczosEvolution."""
    from cyten import krylov_based, sparse
    from cyten_b200.krylov import LanczosEvolution
    sym, msym, A, B = _setup(name, dtype, chi // 4, 8)
    TH = rc.tdot(A, B, ["vR"], ["vL"])
    rdt = rc.Dtype.complex128 if dtype == "complex128" else rc.Dtype.float64
    rb = rc.get_backend("abelian", "numpy")
    X = rc.SymmetricTensor.from_random_normal(TH.codomain, TH.codomain, backend=rb, dtype=rdt)
    K = 0.2 * (X + X.hc)

    class RefOp(sparse.LinearOperator):
        def matvec(self, t):
            return rc.compose(K, t)

    class Op:
        def __init__(self):
            self.k = from_reference(K, backend, msym)

        def matvec(self, t):
            return ct.compose(self.k, t)

    psir, Nr = krylov_based.LanczosEvolution(RefOp(), TH, {"N_max": 30}).run(delta)
    psi, N = LanczosEvolution(Op(), from_reference(TH, backend, msym), {"N_max": 30}).run(delta)
    assert N == Nr
    ref = from_reference(psir, backend, msym)
    diff = ct.norm(ct.linear_combination(1.0, psi, -1.0, ref))
    assert diff <= 1e-10 * max(1.0, ct.norm(ref)), diff


def _ref_bi(T, ncol):
    return np.asarray(T.data.block_inds.to_numpy()).reshape(len(T.data.blocks), ncol)


@pytest.mark.parametrize("name,dtype,chi", CASES)
def test_partial_compose_outer_lq(backend, name, dtype, chi):
    """The remaining members of SURVEY rows a6 / a9: partial_compose (abelian.cpp:2853-2951), the product without
    contracted legs (_compose_no_contraction, :1518-1565, through outer) and lq (:2304-2382)."""
    sym, msym, A, B = _setup(name, dtype, chi // 3, 9)
    TH = rc.tdot(A, B, ["vR"], ["vL"])                       # [vL, p0 | p1, vR]
    th = from_reference(TH, backend, msym)
    rdt = rc.Dtype.complex128 if dtype == "complex128" else rc.Dtype.float64
    rb = rc.get_backend("abelian", "numpy")
    p = TH.codomain[1]
    Op = rc.SymmetricTensor.from_random_normal([p], [p], backend=rb, dtype=rdt, labels=["q", "q*"])
    R, r = rc.tensors.partial_compose(TH, Op, 1), ct.partial_compose(th, from_reference(Op, backend, msym), 1)
    assert list(R.labels) == r.labels
    _same(r, R, tol=1e-13)                                   # an operator applied to a codomain leg
    dleg = TH.domain[1]
    Op2 = rc.SymmetricTensor.from_random_normal([dleg], [dleg], backend=rb, dtype=rdt, labels=["r", "r*"])
    first = TH.num_legs - 2                                  # leg index of domain factor 1
    R, r = rc.tensors.partial_compose(TH, Op2, first), ct.partial_compose(th, from_reference(Op2, backend, msym), first)
    assert list(R.labels) == r.labels
    _same(r, R, tol=1e-13)                                   # ... and to a domain leg
    with pytest.raises(ValueError):                          # the full codomain is compose's job
        full = rc.SymmetricTensor.from_random_normal(TH.codomain, TH.codomain, backend=rb, dtype=rdt)
        ct.partial_compose(th, from_reference(full, backend, msym), 0)
    Bl = rc.SymmetricTensor.from_random_normal([p], [p], backend=rb, dtype=rdt, labels=["x", "y"])
    O, o = rc.outer(A, Bl), ct.outer(from_reference(A, backend, msym), from_reference(Bl, backend, msym))
    assert list(O.labels) == o.labels
    _same(o, O, tol=1e-14)                                   # no contracted leg: every pair of blocks
    Lr, Qr = rc.lq(TH)
    l, q = ct.lq(th)
    assert np.array_equal(l.data.block_inds, _ref_bi(Lr, 3)) and np.array_equal(q.data.block_inds, _ref_bi(Qr, 3))
    _same(ct.compose(l, q), TH, tol=1e-12)                   # L Q = theta (the factors: gauge)
    if q.data.num_blocks:
        qq = ct.compose(q, ct.dagger(q)).to_dense()
        n = qq.shape[0]
        assert np.allclose(qq.reshape(n, n), np.eye(n), rtol=0, atol=1e-12)      # Q Q^dagger = 1 on the new leg

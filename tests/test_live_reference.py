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

"""Krylov vector operations and the Lanczos loop built on them, checked against dense numpy
on both backends (numpy oracle of the C-ABI on CPU; CUDA library on the GPU).  The same operations are compared with
the unmodified reference in tests/test_live_reference.py.
"""
from __future__ import annotations

import numpy as np
import pytest

from cyten_b200 import tensors as ct
from cyten_b200.krylov import GraphedOperator, LanczosGroundState, lanczos
from cyten_b200.workloads import HEFF_LABELS, heff_spaces, random_tensor


def _theta(be, chi, dtype, seed, drop=None):
    cod, dom = heff_spaces(chi, 2, 1.0)["theta"]
    t = random_tensor(cod, dom, be, seed=seed, dtype=dtype, labels=HEFF_LABELS["theta"])
    if drop is not None:                       # a tensor with fewer blocks: linear_combination must take the union
        blocks, bi = t.to_numpy_blocks(), t.data.block_inds
        keep = [i for i in range(len(blocks)) if i % drop]
        t = ct.SymmetricTensor.from_numpy_blocks([blocks[i] for i in keep], bi[keep], t.codomain, t.domain, be,
                                                 labels=t.labels, dtype=dtype)
    return t


@pytest.mark.parametrize("dtype", ["float64", "complex128"])
def test_vector_ops_against_dense(backend, dtype):
    a, b, c = _theta(backend, 48, dtype, 1), _theta(backend, 48, dtype, 2), _theta(backend, 48, dtype, 3, drop=3)
    A, B, Cd = a.to_dense(), b.to_dense(), c.to_dense()
    assert c.data.num_blocks < a.data.num_blocks
    tol = 1e-13 * np.linalg.norm(A) * np.linalg.norm(B)
    assert abs(ct.inner(a, b) - np.vdot(A, B)) <= tol
    assert abs(ct.inner(a, c) - np.vdot(A, Cd)) <= tol and abs(ct.inner(c, a) - np.vdot(Cd, A)) <= tol
    assert abs(ct.norm(a) - np.linalg.norm(A)) <= 1e-13 * np.linalg.norm(A)
    co = (0.3 - 0.2j, 1.7 + 0.1j) if dtype == "complex128" else (0.3, -1.7)
    assert np.allclose(ct.linear_combination(co[0], a, co[1], b).to_dense(), co[0] * A + co[1] * B, rtol=0, atol=1e-14 * 10)
    for x, X, y, Y in ((a, A, c, Cd), (c, Cd, a, A)):
        r = ct.linear_combination(co[0], x, co[1], y)
        assert r.data.num_blocks == a.data.num_blocks
        assert np.array_equal(r.data.block_inds, a.data.block_inds)         # union, lexsorted
        assert np.allclose(r.to_dense(), co[0] * X + co[1] * Y, rtol=0, atol=1e-14 * 10)
    assert np.allclose((a - b).to_dense(), A - B, rtol=0, atol=1e-14 * 10)
    assert np.allclose((2.5 * a).to_dense(), 2.5 * A) and np.allclose((a / 4).to_dense(), A / 4) and np.allclose((-a).to_dense(), -A)
    assert (0.0 * a).data.num_blocks == 0
    if dtype == "float64":
        z = (1j * a)
        assert z.dtype == "complex128" and np.allclose(z.to_dense(), 1j * A)
    # dagger: all axes reversed + conjugation; legs swap roles
    d = ct.dagger(a)
    assert d.labels == ["vR*", "p1*", "p0*", "vL*"]
    assert d.codomain == a.domain and d.domain == a.codomain
    assert np.array_equal(d.to_dense(), np.conj(A).transpose(3, 2, 1, 0))
    assert abs(ct.inner(d, b, do_dagger=False) - np.vdot(A, B)) <= tol        # trace(a^dagger b)
    assert np.array_equal(ct.dagger(d).to_dense(), A)


@pytest.mark.parametrize("dtype", ["float64", "complex128"])
def test_scale_axis_is_compose_with_the_diagonal(backend, dtype):
    th = _theta(backend, 40, dtype, 4)
    u, s, vh = ct.svd(th)
    sd = s.to_numpy()
    U, Vh = u.to_dense(), vh.to_dense()
    us, svh = ct.scale_axis(u, s, -1), ct.scale_axis(vh, s, 0)
    assert np.allclose(us.to_dense(), U * sd[None, None, :], rtol=0, atol=1e-14 * sd.max())
    assert np.allclose(svh.to_dense(), Vh * sd[:, None, None], rtol=0, atol=1e-14 * sd.max())
    rec = ct.compose(us, vh).to_dense()
    assert np.allclose(rec, th.to_dense(), rtol=0, atol=1e-12 * sd.max())
    mid = ct.permute_legs(u, [0, 2], [1])                  # scaled leg in the middle of the block axes
    assert np.allclose(ct.scale_axis(mid, s, 1).to_dense(), mid.to_dense() * sd[None, :, None], rtol=0, atol=1e-14 * sd.max())


class _TwoSided:
    """theta -> K theta + theta K2 with Hermitian K, K2: a Hermitian operator on the two-site space."""

    def __init__(self, K, K2):
        self.K, self.K2 = K, K2

    def matvec(self, t):
        return ct.compose(self.K, t) + ct.compose(t, self.K2)


def _problem(be, chi, dtype):
    th = _theta(be, chi, dtype, 5)
    X = random_tensor(th.codomain.factors, th.codomain.factors, be, seed=6, dtype=dtype, labels=["vL", "p0", "p0*", "vL*"])
    X2 = random_tensor(th.domain.factors, th.domain.factors, be, seed=7, dtype=dtype, labels=["vR*", "p1*", "p1", "vR"])
    K, K2 = X + ct.dagger(X), X2 + ct.dagger(X2)
    return th, _TwoSided(K, K2), K, K2


@pytest.mark.parametrize("dtype", ["float64", "complex128"])
@pytest.mark.parametrize("opts", [{"N_max": 40}, {"N_max": 40, "N_cache": 3}, {"N_max": 40, "reortho": True},
                                  {"N_max": 40, "E_shift": -3.0}])
def test_lanczos_ground_state(backend, dtype, opts):
    th, H, K, K2 = _problem(backend, 16, dtype)
    lz = LanczosGroundState(H, th, dict(opts, P_tol=1e-22))
    E0, psi, N = lz.run()
    assert abs(ct.norm(psi) - 1.0) < 5e-12
    Hpsi = H.matvec(psi)
    residual = ct.norm(Hpsi - E0 * psi)
    assert residual < 1e-6 * (abs(E0) + 1.0), residual
    assert abs(np.real(ct.inner(psi, Hpsi)) - E0) < 1e-10 * (abs(E0) + 1.0)
    assert lz.host_syncs == N + 2                         # one read-back per iteration + the two norms around the loop
    # the Ritz value is an upper bound of the true minimum over the charge sector; compare with a dense solve in the
    # Krylov space enlarged by restarting from psi
    E1, psi1, N1 = lanczos(H, psi, {"N_max": 40, "P_tol": 1e-22})
    assert E1 <= E0 + 1e-10 and abs(E1 - E0) < 1e-8 * (abs(E0) + 1.0)


def test_lanczos_option_errors(backend):
    th, H, _, _ = _problem(backend, 8, "float64")
    with pytest.raises(ValueError):
        LanczosGroundState(H, th, {"N_min": 1})
    with pytest.raises(ValueError):
        LanczosGroundState(H, th, {"N_cache": 1})
    with pytest.raises(ValueError):
        LanczosGroundState(H, th, {"bogus": 1})
    with pytest.raises(ValueError):
        lanczos(H, 0.0 * th + 1e-300 * th, {})            # norm of psi0 too small


@pytest.mark.gpu
def test_lanczos_graph_replay_matches_eager(cuda_backend):
    th, H, _, _ = _problem(cuda_backend, 64, "float64")
    E0, psi, N = lanczos(H, th, {"N_max": 12})
    E1, psi1, N1 = lanczos(GraphedOperator(H), th, {"N_max": 12})
    assert N == N1 and E0 == E1
    assert np.array_equal(psi.to_dense(), psi1.to_dense())


@pytest.mark.parametrize("dtype", ["float64", "complex128"])
@pytest.mark.parametrize("delta", [-0.2j, -0.15, 0.05 - 0.1j])
def test_lanczos_evolution_against_dense_expm(backend, dtype, delta):
    from scipy.linalg import expm
    from cyten_b200.krylov import LanczosEvolution
    th, _, K, _ = _problem(backend, 12, dtype)
    K = 0.1 * K

    class Op:
        def matvec(self, t):
            return ct.compose(K, t)

    psi, N = LanczosEvolution(Op(), th, {"N_max": 40, "P_tol": 1e-15}).run(delta, normalize=False)
    Kd, T = K.to_dense(), th.to_dense()
    n = Kd.shape[0] * Kd.shape[1]
    M = Kd.transpose(0, 1, 3, 2).reshape(n, n)                 # block axes are [vL, p0, p0', vL']
    want = (expm(delta * M) @ T.reshape(n, -1)).reshape(T.shape)
    assert N < 40
    assert psi.dtype == ("float64" if (dtype == "float64" and np.imag(delta) == 0) else "complex128")
    assert np.allclose(psi.to_dense(), want, rtol=0, atol=1e-10 * np.abs(want).max())
    unit, _ = LanczosEvolution(Op(), th, {"N_max": 40, "P_tol": 1e-15}).run(delta, normalize=True)
    assert abs(ct.norm(unit) - 1.0) < 1e-12


def test_outer_partial_compose_lq_against_dense(backend):
    """Self-contained versions of the live-reference checks (dense numpy)."""
    from cyten_b200.spaces import TensorProduct
    th = _theta(backend, 24, "float64", 4)
    A, T = None, th.to_dense()
    p0 = th.codomain.factors[1]
    op = random_tensor([p0], [p0], backend, seed=8, labels=["q", "q*"])
    r = ct.partial_compose(th, op, 1)
    assert r.labels == ["vL", "q", "p1", "vR"]
    assert np.allclose(r.to_dense(), np.einsum("qp,apcd->aqcd", op.to_dense(), T), rtol=0, atol=1e-13)
    dl = th.domain.factors[1]                                  # legs order [vL, p0, p1, vR]: domain = [vR, p1]
    op2 = random_tensor([dl], [dl], backend, seed=9, labels=["r", "r*"])
    r2 = ct.partial_compose(th, op2, 2)
    assert np.allclose(r2.to_dense(), np.einsum("abcd,cr->abrd", T, op2.to_dense()), rtol=0, atol=1e-13)
    o = ct.outer(op, op2)
    assert o.num_codomain_legs == 2 and o.num_domain_legs == 2
    assert np.allclose(o.to_dense(), np.einsum("ab,cd->acdb", op.to_dense(), op2.to_dense()), rtol=0, atol=1e-14)
    l, q = ct.lq(th)
    assert np.allclose(ct.compose(l, q).to_dense(), T, rtol=0, atol=1e-12 * np.abs(T).max())
    qq = ct.compose(q, ct.dagger(q)).to_dense()
    assert np.allclose(qq, np.eye(qq.shape[0]), rtol=0, atol=1e-12)
    empty = TensorProduct([], symmetry=th.symmetry)
    a = random_tensor(th.codomain.factors, empty, backend, seed=3)
    b = random_tensor(empty, th.codomain.factors, backend, seed=4)
    assert np.allclose(ct.compose(a, b).to_dense(), np.multiply.outer(a.to_dense(), b.to_dense()), rtol=0, atol=1e-14)

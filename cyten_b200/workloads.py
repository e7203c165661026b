"""Synthetic workloads of the benchmark configurations.

Deterministic U(1) legs: bond leg with charges -Q..Q and Gaussian multiplicities
``m_q = max(1, floor(chi*g_q/sum g))``, ``g_q = exp(-q^2 / 2 sigma^2)``, remainder added to q = 0;
physical leg charges {-1,+1}; MPO leg D = 5 as charges {-2:1, 0:3, +2:1}.  All blocks present.
"""
from __future__ import annotations

import numpy as np

from .spaces import AbelianSymmetry, ElementarySpace, TensorProduct, u1_symmetry
from .tensors import SymmetricTensor

CONFIG_LEGS = {128: (3, 1.2), 1024: (6, 2.2), 4096: (9, 3.2), 16384: (12, 4.4)}


def u1_bond_leg(chi: int, Q: int, sigma: float, is_dual: bool = False) -> ElementarySpace:
    q = np.arange(-Q, Q + 1)
    g = np.exp(-q.astype(float) ** 2 / (2 * sigma ** 2))
    m = np.maximum(1, np.floor(chi * g / g.sum()).astype(np.int64))
    m[Q] += chi - int(m.sum())
    if m[Q] < 1:
        raise ValueError("chi too small for this (Q, sigma)")
    return ElementarySpace(u1_symmetry, q[:, None], m, is_dual)


def u1_phys_leg(is_dual: bool = False) -> ElementarySpace:
    return ElementarySpace(u1_symmetry, [[-1], [1]], [1, 1], is_dual)


def u1_mpo_leg(is_dual: bool = False) -> ElementarySpace:
    return ElementarySpace(u1_symmetry, [[-2], [0], [2]], [1, 3, 1], is_dual)


# ---- Fermi-Hubbard legs, U(1) x U(1) = (charge N, 2 Sz): BASELINE config 4 / SURVEY section 8(d) ----
u1u1_symmetry = AbelianSymmetry((0, 0))


def hubbard_bond_leg(chi: int, Q: int = 6, sigma=(2.0, 2.0), is_dual: bool = False) -> ElementarySpace:
    """2-D Gaussian multiplicities over (N, 2Sz) in [-Q, Q]^2 with N + 2Sz even (N fermions have 2Sz of
    the same parity), ``m = max(1, floor(chi g / sum g))``, remainder added to (0, 0): many small
    irregular blocks."""
    qs = [(n, s) for n in range(-Q, Q + 1) for s in range(-Q, Q + 1) if (n + s) % 2 == 0]
    q = np.asarray(qs, dtype=np.int64)
    g = np.exp(-q[:, 0].astype(float) ** 2 / (2 * sigma[0] ** 2) - q[:, 1].astype(float) ** 2 / (2 * sigma[1] ** 2))
    m = np.maximum(1, np.floor(chi * g / g.sum()).astype(np.int64))
    i0 = qs.index((0, 0))
    m[i0] += chi - int(m.sum())
    if m[i0] < 1:
        raise ValueError("chi too small for this (Q, sigma)")
    return ElementarySpace.from_defining_sectors(u1u1_symmetry, q, m, is_dual)


def hubbard_phys_leg(is_dual: bool = False) -> ElementarySpace:
    """d = 4: (N, 2Sz) in {(0,0), (1,1), (1,-1), (2,0)}."""
    return ElementarySpace.from_defining_sectors(u1u1_symmetry, [[0, 0], [1, -1], [1, 1], [2, 0]], [1, 1, 1, 1], is_dual)


def random_tensor(codomain, domain, backend, seed=0, dtype="float64", labels=None) -> SymmetricTensor:
    rng = np.random.default_rng(seed)

    def func(shape):
        x = rng.standard_normal(shape)
        if dtype == "complex128":
            x = x + 1j * rng.standard_normal(shape)
        return x
    return SymmetricTensor.from_block_func(func, codomain, domain, backend, labels=labels, dtype=dtype)


def mps_tensor(vL, p, vR, backend, seed=0, dtype="float64") -> SymmetricTensor:
    """``B[vL, p | vR]``: codomain (vL, p), domain (vR)."""
    return random_tensor([vL, p], [vR], backend, seed, dtype, labels=["vL", "p", "vR"])


def compose_flops(plan, dtype="float64") -> float:
    """Algorithmic flops of a compose: ``sum c*M*K*N`` over the reference's matrix_dot list
    (abelian.cpp:1424-1446), c = 2 real / 8 complex."""
    return (8.0 if dtype == "complex128" else 2.0) * float(plan["mkn"])


# ---- host-side block generation (shared by the GPU arm, the oracle arm and the reference arm) ----
def random_blocks(codomain, domain, seed=0, dtype="float64"):
    """(block_inds, blocks) of a tensor with all symmetry-allowed blocks, N(0,1) entries."""
    from .backend import conventional_leg_order
    from .tensors import valid_block_inds
    codomain = codomain if isinstance(codomain, TensorProduct) else TensorProduct(list(codomain))
    if not isinstance(domain, TensorProduct):
        domain = TensorProduct(list(domain), symmetry=codomain.symmetry)
    bi = valid_block_inds(codomain, domain)
    legs = conventional_leg_order(codomain, domain)
    rng = np.random.default_rng(seed)
    blocks = []
    for row in bi:
        shape = tuple(int(l.multiplicities[b]) for l, b in zip(legs, row))
        x = rng.standard_normal(shape)
        if dtype == "complex128":
            x = x + 1j * rng.standard_normal(shape)
        blocks.append(x)
    return bi, blocks


def heff_spaces(chi: int, Q: int | None = None, sigma: float | None = None):
    """Legs of the two-site effective Hamiltonian (toycodes/tenpy_toycodes/d_dmrg.py:40-88) in the
    already-bent form HEffective.__init__ produces:
      theta: [v, p] <- [v, p*]      LP: [v, w*] <- [v]      W1 = W2: [p, w*] <- [w*, p]      RP: [v, w] <- [v]
    """
    if Q is None:
        Q, sigma = CONFIG_LEGS[chi]
    v, p, w = u1_bond_leg(chi, Q, sigma), u1_phys_leg(), u1_mpo_leg()
    return dict(theta=([v, p], [v, p.dual]), LP=([v, w.dual], [v]), W1=([p, w.dual], [w.dual, p]),
                W2=([p, w.dual], [w.dual, p]), RP=([v, w], [v]))


HEFF_LABELS = dict(theta=["vL", "p0", "p1", "vR"], LP=["vL", "wL*", "vL*"], W1=["p0", "wC", "p0*", "wL"],
                   W2=["p1", "wR", "p1*", "wC"], RP=["vR*", "wR*", "vR"])
HEFF_SEEDS = dict(theta=1, LP=2, W1=3, W2=4, RP=5)


class HEffective:
    """Two-site DMRG effective Hamiltonian; ``matvec`` is the call sequence of
    ``HEffective.matvec`` (d_dmrg.py:70-88): 5 x permute_legs interleaved with 4 x compose."""

    def __init__(self, LP, RP, W1, W2, ops):
        self.LP, self.RP, self.W1, self.W2, self.ct = LP, RP, W1, W2, ops

    def matvec(self, theta):
        ct = self.ct
        x = ct.permute_legs(theta, ["vL"], ["vR", "p1", "p0"])
        x = ct.compose(self.LP, x)                                  # vL wL* p0 p1 vR
        x = ct.permute_legs(x, ["wL*", "p0"], ["vL", "vR", "p1"])
        x = ct.compose(self.W1, x)                                  # p0 wC p1 vR vL
        x = ct.permute_legs(x, ["wC", "p1"], ["p0", "vL", "vR"])
        x = ct.compose(self.W2, x)                                  # p1 wR vR vL p0
        x = ct.permute_legs(x, ["vL", "p0", "p1"], ["vR", "wR"])
        x = ct.compose(x, self.RP)                                  # vL p0 p1 vR
        x = ct.permute_legs(x, domain=["vR", "p1"])
        return x


def build_heff(chi: int, backend, dtype="float64", Q=None, sigma=None):
    """(HEffective, theta) on ``backend`` plus the host blocks of theta (for the end-to-end arm)."""
    from . import tensors as ops
    sp = heff_spaces(chi, Q, sigma)
    T, host = {}, {}
    for name, (cod, dom) in sp.items():
        bi, blocks = random_blocks(cod, dom, HEFF_SEEDS[name], dtype)
        host[name] = (bi, blocks)
        T[name] = SymmetricTensor.from_numpy_blocks(blocks, bi, cod, dom, backend, labels=HEFF_LABELS[name], dtype=dtype)
    return HEffective(T["LP"], T["RP"], T["W1"], T["W2"], ops), T["theta"], host, sp

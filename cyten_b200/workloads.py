"""Synthetic workloads of the benchmark configurations.

Deterministic U(1) legs: bond leg with charges -Q..Q and Gaussian multiplicities
``m_q = max(1, floor(chi*g_q/sum g))``, ``g_q = exp(-q^2 / 2 sigma^2)``, remainder added to q = 0;
physical leg charges {-1,+1}; MPO leg D = 5 as charges {-2:1, 0:3, +2:1}.  All blocks present.
"""
from __future__ import annotations

import numpy as np

from .spaces import ElementarySpace, TensorProduct, u1_symmetry
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

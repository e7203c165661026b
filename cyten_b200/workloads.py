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


# ---- BASELINE config 5: one bond update of a two-site DMRG sweep ---------------------------------------------------------
class NativeOps:
    """The operations ``SweepStep`` needs, on cyten_b200 tensors (the same interface is implemented over the reference's
    ``cyten`` module by ``bench.py`` / the tests, so that ONE call sequence runs on both)."""

    def __init__(self):
        from . import tensors as t
        self.t = t

    @staticmethod
    def relabel(T, mapping):
        labels = [mapping.get(l, l) for l in T.labels]
        if hasattr(T, "codomain"):
            return SymmetricTensor(T.data, T.codomain, T.domain, T.backend, labels)
        from .tensors import DiagonalTensor
        return DiagonalTensor(T.data, T.leg, T.backend, labels)

    def permute_legs(self, T, codomain=None, domain=None):
        return self.t.permute_legs(T, codomain, domain)

    def compose(self, A, B, relabel1=None, relabel2=None):
        if relabel1:
            A = self.relabel(A, relabel1)
        if relabel2:
            B = self.relabel(B, relabel2)
        return self.t.compose(A, B)

    def truncated_svd(self, theta, new_labels, chi_max, svd_min):
        U, S, Vh, err, _ = self.t.truncated_svd(theta, new_labels=(new_labels[0], new_labels[0], new_labels[1], new_labels[1]),
                                                chi_max=chi_max, svd_min=svd_min)
        return U, S, Vh, err

    def scale_axis(self, T, D, leg):
        return self.t.scale_axis(T, D, leg)

    def pinv(self, D, cutoff):
        return self.t.pinv(D, cutoff)

    def hc(self, T):
        return self.t.dagger(T)

    def norm(self, T):
        return float(self.t.norm(T))

    def scale(self, T, a):
        return self.t.scalar_multiply(a, T)

    # multi-GPU hooks (cyten_b200.sharding.ShardedSweepOps): identity on one GPU
    @staticmethod
    def shard(T, leg):
        return T

    @staticmethod
    def gather(T):
        return T


class SweepStep:
    """``DMRGEngine.update_bond`` of the reference's toy code (toycodes/tenpy_toycodes/d_dmrg.py:217-274, with
    ``split_truncate_theta``, toycodes/tenpy_toycodes/a_mps.py:217-250), call for call, on synthetic tensors:

    1. bend the environments / MPO tensors into the effective Hamiltonian (4 permute_legs),
    2. ``nmv`` applications of H_eff to theta (the reference runs Lanczos here; a fixed number of matvecs + a
       normalisation is the same kernel work without the data-dependent iteration count),
    3. ``truncated_svd(theta, chi_max, svd_min=eps)`` + the permutation of ``B``,
    4. ``scale_axis`` with ``pinv`` of the Schmidt values (the new B matrices),
    5. ``update_LP`` (2 scale_axis, hc, 4 permute_legs, 3 compose) and ``update_RP`` (hc, 5 permute_legs, 3 compose).

    ``ops``: :class:`NativeOps` or the same interface over the reference.  ``timings``: optional dict that receives the
    seconds per part when ``sync`` (a callable that waits for the device) is given."""

    def __init__(self, ops, LP, RP, W, S_left, chi_max, eps=1e-14, nmv=1, sync=None, fused=False):
        self.ops, self.LP, self.RP, self.W, self.S_left = ops, LP, RP, W, S_left
        self.chi_max, self.eps, self.nmv, self.sync = chi_max, eps, nmv, sync
        self.fused = fused          # native tensors only: H_eff as GEMM / one fused middle pass / GEMM (cyten_b200/fused.py)
        self.timings = {}

    def _mark(self, name, t0):
        import time
        if self.sync is not None:
            self.sync()
            self.timings[name] = self.timings.get(name, 0.0) + time.perf_counter() - t0
        return time.perf_counter()

    def heff(self):
        o = self.ops
        LP = o.permute_legs(self.LP, ["vL", "wL*"], ["vL*"])
        RP = o.permute_legs(self.RP, ["vR*", "wR*"], ["vR"])
        W1 = o.relabel(o.permute_legs(self.W, ["p", "wR"], ["wL", "p*"]), {"p": "p0", "p*": "p0*", "wR": "wC"})
        W2 = o.relabel(o.permute_legs(self.W, ["p", "wR"], ["wL", "p*"]), {"p": "p1", "p*": "p1*", "wL": "wC"})
        H = HEffective(LP, RP, W1, W2, o)
        if self.fused:
            from .fused import FusedHEffective
            return FusedHEffective(H)
        return H

    def update_bond(self, theta0):
        import time
        o = self.ops
        t0 = time.perf_counter()
        H = self.heff()
        theta = theta0
        for _ in range(self.nmv):
            theta = H.matvec(theta)
        theta = o.scale(theta, 1.0 / o.norm(theta))
        t0 = self._mark("matvec", t0)
        A, S, B, err = o.truncated_svd(theta, ["vR", "vL"], self.chi_max, self.eps)
        B = o.permute_legs(B, codomain=["vL", "p1"])
        A, B = o.relabel(A, {"p0": "p"}), o.relabel(B, {"p1": "p"})
        t0 = self._mark("truncated_svd", t0)
        G = o.scale_axis(A, o.pinv(self.S_left, self.eps), "vL")
        Bi = o.scale_axis(G, S, "vR")
        t0 = self._mark("scale_axis", t0)
        LP_new = self.update_LP(Bi, S)
        t0 = self._mark("update_LP", t0)
        RP_new = self.update_RP(B)
        self._mark("update_RP", t0)
        return dict(Bi=Bi, Bj=B, S=S, err=err, LP=LP_new, RP=RP_new, theta=theta)

    def update_LP(self, B, S_right):
        o = self.ops
        G = o.scale_axis(B, o.pinv(S_right, self.eps), "vR")
        A = o.scale_axis(G, self.S_left, "vL")
        Ac = o.permute_legs(o.hc(A), codomain=["vR*", "p*"])
        Ac = o.shard(Ac, "vR*")       # N ranks: vR* stays open to the end, so every rank builds disjoint blocks of the result
        LP = o.compose(Ac, self.LP)                                              # vR* p* wL* vL*
        LP = o.permute_legs(LP, ["vL*", "vR*"], ["wL*", "p*"])
        LP = o.compose(LP, self.W)                                               # vL* vR* wR p*
        LP = o.permute_legs(LP, ["vR*", "wR"], ["vL*", "p*"])
        LP = o.compose(LP, A, relabel1={"vR*": "vL", "wR": "wL*"}, relabel2={"vR": "vL*"})
        return o.gather(o.permute_legs(LP, domain=["vL*", "wL*"]))

    def update_RP(self, B):
        o = self.ops
        Bc = o.permute_legs(o.hc(B), ["p*", "vL*"], ["vR*"])
        Bc = o.shard(Bc, "vL*")       # N ranks: vL* (the new vR) stays open to the end
        RP = o.compose(Bc, self.RP)                                              # p* vL* vR* wR*
        RP = o.permute_legs(RP, ["vL*", "vR*"], ["p*", "wR*"])
        W_ = o.permute_legs(self.W, ["p", "wR"], ["wL", "p*"])
        RP = o.compose(RP, W_)                                                   # vL* vR* p* wL
        RP = o.permute_legs(RP, ["wL", "vL*"], ["p*", "vR*"])
        B_ = o.permute_legs(B, ["p", "vR"], ["vL"])
        RP = o.compose(RP, B_, relabel1={"vL*": "vR", "wL": "wR*"}, relabel2={"vL": "vR*"})
        return o.gather(o.permute_legs(RP, ["vR"], ["wR*", "vR*"]))


def build_sweep_step(chi: int, backend, nmv=1, chi_max=None, eps=1e-14, Q=None, sigma=None, sync=None, fused=False):
    """(SweepStep, theta0) on ``backend``: the H_eff tensors of :func:`build_heff` un-bent into the toy code's storage
    form (``LP[vL | vL*, wL*]``, ``RP[vR | wR*, vR*]``, ``W[wL, p | p*, wR]``), positive Schmidt values on the left bond."""
    from . import tensors as ops
    from .tensors import DiagonalTensor
    H, theta, host, sp = build_heff(chi, backend, Q=Q, sigma=sigma)
    o = NativeOps()
    LP = ops.permute_legs(H.LP, ["vL"], ["vL*", "wL*"])
    RP = ops.permute_legs(H.RP, ["vR"], ["wR*", "vR*"])
    W = ops.permute_legs(o.relabel(H.W1, {"p0": "p", "p0*": "p*", "wC": "wR"}), ["wL", "p"], ["p*", "wR"])
    v = sp["theta"][0][0]
    S_left = diagonal_from_numpy(schmidt_values(v), v, backend, ["vL", "vR"])
    return SweepStep(o, LP, RP, W, S_left, chi_max or chi, eps, nmv, sync, fused), theta


def schmidt_values(leg, seed=6) -> np.ndarray:
    """Positive, normalised synthetic Schmidt values over the basis of ``leg`` (internal order)."""
    rng = np.random.default_rng(seed)
    s = np.abs(rng.standard_normal(int(leg.dim))) + 0.05
    return s / np.linalg.norm(s)


def diagonal_from_numpy(values: np.ndarray, leg, backend, labels=None):
    """DiagonalTensor over ``leg`` from the dense diagonal (internal basis order), one block per sector."""
    import torch
    from .data import AbelianBackendData
    from .tensors import DiagonalTensor
    ns = leg.num_sectors
    bi = np.stack([np.arange(ns), np.arange(ns)], axis=1).astype(np.int64)
    data = AbelianBackendData.empty("float64", backend.device, bi, np.asarray(leg.multiplicities, np.int64).reshape(-1, 1))
    host = np.zeros(data.buf.numel())
    for i, off in enumerate(data.offsets):
        b0, b1 = leg.slices[i]
        host[off:off + (b1 - b0)] = values[b0:b1]
    data.buf.copy_(torch.from_numpy(host))
    return DiagonalTensor(data, leg, backend, labels or [None, None])

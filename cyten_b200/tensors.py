"""Tensor-level API mirroring the reference's free functions for the hot path.

``tdot`` = 2 x ``permute_legs`` + ``compose`` (src/tensors/ops_algebra.cpp:1908-1917);
``svd`` / ``qr`` / ``eigh`` = ``_decomposition_prepare`` (combine_legs) + backend call + ``split_legs``
(src/tensors/helpers.cpp:596-647, src/tensors/decompositions.cpp:314-424, 466-508, 554-617);
``truncated_svd`` (decompositions.cpp:672-709).  Leg bookkeeping follows the reference:
``tensor.legs = [*codomain, *(l.dual for l in reversed(domain))]`` and every block has the axis order
``[*codomain, *reversed(domain)]``.
"""
from __future__ import annotations

from typing import Sequence

import numpy as np

from .backend import B200AbelianBackend, conventional_leg_order
from .data import AbelianBackendData
from .spaces import AbelianLegPipe, ElementarySpace, TensorProduct

__all__ = ["SymmetricTensor", "DiagonalTensor", "permute_legs", "combine_legs", "split_legs",
           "compose", "tdot", "svd", "truncated_svd", "qr", "eigh", "norm", "inner", "linear_combination",
           "scalar_multiply", "dagger", "scale_axis", "outer", "partial_compose", "lq", "pinv"]


class SymmetricTensor:
    """Carrier of (data, codomain, domain, backend), like cyten's SymmetricTensor
    (include/cyten/tensors/symmetric_tensor.h:34-42)."""

    def __init__(self, data: AbelianBackendData, codomain, domain, backend: B200AbelianBackend,
                 labels=None):
        self.data = data
        self.codomain = codomain if isinstance(codomain, TensorProduct) else TensorProduct(list(codomain))
        if not isinstance(domain, TensorProduct):
            domain = TensorProduct(list(domain), symmetry=self.codomain.symmetry)
        self.domain = domain
        self.backend = backend
        self.num_codomain_legs = self.codomain.num_factors
        self.num_domain_legs = self.domain.num_factors
        self.num_legs = self.num_codomain_legs + self.num_domain_legs
        self.labels = list(labels) if labels is not None else [None] * self.num_legs
        if data.block_inds.shape[1] != self.num_legs:
            raise ValueError("block_inds do not match the number of legs")

    @property
    def symmetry(self):
        return self.codomain.symmetry

    @property
    def dtype(self):
        return self.data.dtype

    @property
    def legs(self):
        return [*self.codomain.factors, *(l.dual for l in reversed(self.domain.factors))]

    def get_leg_co_domain(self, i):
        return conventional_leg_order(self.codomain, self.domain)[i]

    def get_leg_idcs(self, which) -> list[int]:
        res = []
        for w in (which if isinstance(which, (list, tuple)) else [which]):
            if isinstance(w, str):
                res.append(self.labels.index(w))
            else:
                res.append(int(w) % self.num_legs)
        return res

    # -- construction -------------------------------------------------------------------------------
    @classmethod
    def from_numpy_blocks(cls, blocks, block_inds, codomain, domain, backend, labels=None, dtype=None):
        """From host blocks in the reference's layout (e.g. ``T.data.blocks`` / ``T.data.block_inds``)."""
        data = AbelianBackendData.from_numpy_blocks(blocks, block_inds, backend.device, dtype=dtype)
        return cls(data, codomain, domain, backend, labels)

    @classmethod
    def from_block_func(cls, func, codomain, domain, backend, labels=None, dtype="float64"):
        """All symmetry-allowed blocks filled by ``func(shape) -> ndarray`` (valid_block_inds,
        abelian.cpp:329-399: fused codomain charge == fused domain charge)."""
        codomain = codomain if isinstance(codomain, TensorProduct) else TensorProduct(list(codomain))
        if not isinstance(domain, TensorProduct):
            domain = TensorProduct(list(domain), symmetry=codomain.symmetry)
        bi = valid_block_inds(codomain, domain)
        legs = conventional_leg_order(codomain, domain)
        blocks = [func(tuple(int(l.multiplicities[b]) for l, b in zip(legs, row))) for row in bi]
        return cls.from_numpy_blocks(blocks, bi, codomain, domain, backend, labels, dtype=dtype)

    def to_numpy_blocks(self):
        return self.data.to_numpy_blocks()

    def to_dense(self) -> np.ndarray:
        """Dense array in the internal (sector-sorted) basis of every leg, axis order
        ``[*codomain, *reversed(domain)]`` (to_dense_block, abelian.cpp:3570-3590, without basis_perm)."""
        legs = conventional_leg_order(self.codomain, self.domain)
        out = np.zeros([l.dim for l in legs], dtype=self.data.dtype)
        for row, blk in zip(self.data.block_inds, self.data.to_numpy_blocks()):
            out[tuple(slice(*l.slices[b]) for l, b in zip(legs, row))] = blk
        return out


class DiagonalTensor:
    """Diagonal tensor on one leg (singular values / eigenvalues): blocks are vectors, block_inds
    rows are ``[i, i]``."""

    def __init__(self, data: AbelianBackendData, leg: ElementarySpace, backend, labels=None):
        self.data, self.leg, self.backend = data, leg, backend
        self.labels = list(labels) if labels is not None else [None, None]

    def to_numpy(self) -> np.ndarray:
        out = np.zeros(self.leg.dim, dtype=self.data.dtype)
        for row, blk in zip(self.data.block_inds, self.data.to_numpy_blocks()):
            out[slice(*self.leg.slices[row[0]])] = blk
        return out


def valid_block_inds(codomain: TensorProduct, domain: TensorProduct) -> np.ndarray:
    """Lexsorted block_inds of all blocks allowed by the charge rule (abelian.cpp:329-399)."""
    legs = conventional_leg_order(codomain, domain)
    sym = codomain.symmetry
    nsecs = [l.num_sectors for l in legs]
    if not legs:
        return np.zeros((1, 0), dtype=np.int64)
    # Sorted merge instead of the reference's brute force over the full product grid:
    # fuse the codomain grid and the domain grid separately, then join the two lists on the fused charge --
    # |cod grid| + |dom grid| + |result| rows instead of |cod grid| x |dom grid|.
    ncod = codomain.num_factors
    nsym = sym.sector_ind_len

    def side(side_legs):
        if not side_legs:
            return np.zeros((1, 0), dtype=np.int64), np.zeros((1, nsym), dtype=np.int64)
        g = np.indices([l.num_sectors for l in side_legs]).reshape(len(side_legs), -1).T
        q = sym.multiple_fusion_broadcast([l.sector_decomposition[g[:, i]] for i, l in enumerate(side_legs)])
        return g, np.asarray(q, dtype=np.int64).reshape(len(g), nsym)
    gc, qc = side(legs[:ncod])
    gd, qd = side(legs[ncod:])
    both, inv = np.unique(np.concatenate([qc, qd]), axis=0, return_inverse=True)
    inv = np.asarray(inv).reshape(-1)
    kc, kd = inv[:len(qc)], inv[len(qc):]
    oc, od = np.argsort(kc, kind="stable"), np.argsort(kd, kind="stable")
    nc, nd = np.bincount(kc, minlength=len(both)), np.bincount(kd, minlength=len(both))
    sc, sd = np.concatenate([[0], np.cumsum(nc)]), np.concatenate([[0], np.cumsum(nd)])
    parts = []
    for k in np.nonzero((nc > 0) & (nd > 0))[0]:
        rc, rd = oc[sc[k]:sc[k + 1]], od[sd[k]:sd[k + 1]]
        parts.append(np.concatenate([np.repeat(gc[rc], len(rd), axis=0), np.tile(gd[rd], (len(rc), 1))], axis=1))
    bi = (np.concatenate(parts) if parts else np.zeros((0, len(legs)), dtype=np.int64)).astype(np.int64)
    return bi[np.lexsort(bi.T)] if len(bi) else bi


# --------------------------------------------------------------------------------------------------
def permute_legs(T: SymmetricTensor, codomain: Sequence | None = None, domain: Sequence | None = None):
    """Like ``cyten.permute_legs``: the listed legs end up in the new codomain / domain in the given
    order; legs crossing between codomain and domain are dualised (src/tensors/ops_legs.cpp:332-640)."""
    n = T.num_legs
    if codomain is None and domain is None:
        raise ValueError("need codomain or domain")
    cod = T.get_leg_idcs(list(codomain)) if codomain is not None else None
    dom = T.get_leg_idcs(list(domain)) if domain is not None else None
    if cod is None:
        cod = [i for i in range(n) if i not in dom]
    if dom is None:
        dom = [i for i in reversed(range(n)) if i not in cod]
    if cod == list(range(T.num_codomain_legs)) and dom == list(reversed(range(T.num_codomain_legs, n))):
        return T
    legs = T.legs
    new_codomain = TensorProduct([legs[i] for i in cod], symmetry=T.symmetry)
    new_domain = TensorProduct([legs[i].dual for i in dom], symmetry=T.symmetry)
    data = T.backend.permute_legs(T, cod, dom, new_codomain, new_domain)
    return SymmetricTensor(data, new_codomain, new_domain, T.backend,
                           [T.labels[i] for i in cod] + [T.labels[i] for i in reversed(dom)])


def make_pipe(legs, is_dual: bool, in_domain: bool) -> AbelianLegPipe:
    """TensorBackend::make_pipe (tensor_backend.cpp:118-136): ``combine_cstyle = not is_dual``."""
    legs = list(legs)
    key = (tuple(l.sig for l in legs), bool(is_dual))
    pipe = _PIPE_CACHE.get(key)
    if pipe is None:
        if len(_PIPE_CACHE) > 4096:
            _PIPE_CACHE.clear()
        pipe = _PIPE_CACHE[key] = AbelianLegPipe(legs, is_dual=is_dual, combine_cstyle=not is_dual)
    return pipe


# AbelianLegPipe.prepare is pure integer work on the legs' sector tables: pipes are immutable, so one
# instance per (legs, duality) serves every combine_legs of a sweep
_PIPE_CACHE: dict = {}


def combine_legs(T: SymmetricTensor, *groups, pipe_dualities=None):
    """Like ``cyten.combine_legs``: every group is a list of ADJACENT legs lying entirely in the
    codomain or in the domain (ascending leg indices); returns the tensor with pipes."""
    groups = [T.get_leg_idcs(list(g)) for g in groups]
    ncod = T.num_codomain_legs
    pipes = []
    cod_f, dom_rev = list(T.codomain.factors), list(reversed(T.domain.factors))   # by leg index
    new_by_idx = {}
    for gi, g in enumerate(groups):
        if g != list(range(g[0], g[0] + len(g))):
            raise ValueError("legs to combine must be adjacent and ascending")
        in_domain = g[0] >= ncod
        if (g[-1] >= ncod) != in_domain:
            raise ValueError("a group must lie entirely in the codomain or the domain")
        if in_domain:
            # domain factors in domain order = reversed leg order
            factors = [dom_rev[i - ncod] for i in reversed(g)]
        else:
            factors = [cod_f[i] for i in g]
        pd = bool(pipe_dualities[gi]) if pipe_dualities is not None else False
        # the pipe is result.domain[k], which has the opposite duality of result.legs[-1-k]
        # (src/tensors/ops_legs.cpp:885-899)
        pipe = make_pipe(factors, (not pd) if in_domain else pd, in_domain)
        pipes.append(pipe)
        new_by_idx[g[0]] = pipe
    skip = {i for g in groups for i in g[1:]}
    new_cod = [new_by_idx.get(i, cod_f[i] if i < ncod else None) for i in range(ncod) if i not in skip]
    new_dom_rev = [new_by_idx.get(i, dom_rev[i - ncod]) for i in range(ncod, T.num_legs) if i not in skip]
    new_codomain = TensorProduct(new_cod, symmetry=T.symmetry)
    new_domain = TensorProduct(list(reversed(new_dom_rev)), symmetry=T.symmetry)
    data = T.backend.combine_legs(T, groups, pipes, new_codomain, new_domain)
    labels = []
    for i in range(T.num_legs):
        if i in skip:
            continue
        g = next((g for g in groups if g[0] == i), None)
        labels.append(T.labels[i] if g is None else "(" + ".".join(str(T.labels[k]) for k in g) + ")")
    return SymmetricTensor(data, new_codomain, new_domain, T.backend, labels)


def split_legs(T: SymmetricTensor, legs=None):
    """Like ``cyten.split_legs``: splits the given pipes (default: all)."""
    order = conventional_leg_order(T.codomain, T.domain)
    if legs is None:
        idcs = [i for i, l in enumerate(order) if isinstance(l, AbelianLegPipe)]
    else:
        idcs = sorted(T.get_leg_idcs(list(legs)))
    if not idcs:
        return T
    ncod = T.num_codomain_legs
    new_cod, new_dom_rev, labels = [], [], []
    for i, l in enumerate(order):
        tgt = new_cod if i < ncod else new_dom_rev
        if i in idcs:
            tgt.extend(l.legs if i < ncod else reversed(l.legs))
            lab = T.labels[i]
            parts = lab[1:-1].split(".") if isinstance(lab, str) and lab.startswith("(") else [None] * l.num_legs
            labels.extend(parts)
        else:
            tgt.append(l)
            labels.append(T.labels[i])
    new_codomain = TensorProduct(new_cod, symmetry=T.symmetry)
    new_domain = TensorProduct(list(reversed(new_dom_rev)), symmetry=T.symmetry)
    data = T.backend.split_legs(T, idcs, new_codomain, new_domain)
    return SymmetricTensor(data, new_codomain, new_domain, T.backend, labels)


def compose(A: SymmetricTensor, B: SymmetricTensor):
    """``A @ B``: contracts A.domain with B.codomain (helpers.cpp:259-294)."""
    if A.domain != B.codomain:
        raise ValueError("compose: A.domain != B.codomain")
    data = A.backend.compose(A, B)
    return SymmetricTensor(data, A.codomain, B.domain, A.backend,
                           A.labels[:A.num_codomain_legs] + B.labels[B.num_codomain_legs:])


def outer(A: SymmetricTensor, B: SymmetricTensor):
    """Tensor product without contraction (ops_algebra.cpp:1097-1178 -> ``_compose_no_contraction``,
    abelian.cpp:1518-1565): legs ``[*A.legs, *B.legs]``, ``A``'s legs bent into the codomain and ``B``'s into the
    domain for the product, then codomain ``A.codomain + B.codomain`` and domain ``A.domain + B.domain`` (the
    reference's convention)."""
    nA, nB = A.num_legs, B.num_legs
    A2 = permute_legs(A, codomain=list(range(nA)), domain=[])
    B2 = permute_legs(B, codomain=[], domain=list(reversed(range(nB))))
    R = compose(A2, B2)                       # legs [*A.legs, *B.legs]
    cA, cB = A.num_codomain_legs, B.num_codomain_legs
    cod = list(range(cA)) + [nA + i for i in range(cB)]
    dom = [nA - 1 - j for j in range(A.num_domain_legs)] + [nA + nB - 1 - j for j in range(B.num_domain_legs)]
    return permute_legs(R, cod, dom)


def partial_compose(T1: SymmetricTensor, T2: SymmetricTensor, first_leg):
    """Applies ``T2`` to a consecutive part of ``T1``'s codomain (``T2.domain`` is contracted, ``T2.codomain`` takes
    its place) or of its domain (``T2.codomain`` is contracted, ``T2.domain`` takes its place), starting at leg
    ``first_leg`` of ``T1``; ops_algebra.cpp:1180-1370 -> abelian.cpp:2853-2951.  Here: one ``tdot`` (the
    permutations are deferred and folded into the GEMM operand re-layout) and one deferred permutation back."""
    n1, c1, d1 = T1.num_legs, T1.num_codomain_legs, T1.num_domain_legs
    n2, c2, d2 = T2.num_legs, T2.num_codomain_legs, T2.num_domain_legs
    f = T1.get_leg_idcs([first_leg])[0]
    if f < c1:
        num, last = d2, f + d2 - 1
        if num == 0:
            raise ValueError("Use compose or outer when no legs are to be contracted")
        if last >= c1:
            raise ValueError("Not all legs to be contracted are in the (co)domain")
        if num >= c1:
            raise ValueError("Use compose for contracting the full (co)domain")
        if list(T1.codomain.factors[f:last + 1]) != list(T2.domain.factors):
            raise ValueError("partial_compose: incompatible legs")
        R = tdot(T2, T1, [n2 - 1 - j for j in range(num)], [f + j for j in range(num)])   # [T2.codomain, T1 open legs]

        def pos(i):
            return c2 + (i if i < f else i - num)
        cod = [pos(i) for i in range(f)] + list(range(c2)) + [pos(i) for i in range(last + 1, c1)]
        dom = [pos(n1 - 1 - j) for j in range(d1)]
        return permute_legs(R, cod, dom)
    num, last = c2, f + c2 - 1
    if num == 0:
        raise ValueError("Use compose or outer when no legs are to be contracted")
    if last >= n1:
        raise ValueError("Not all legs to be contracted are in the (co)domain")
    if num >= d1:
        raise ValueError("Use compose for contracting the full (co)domain")
    df, dl = n1 - 1 - last, n1 - 1 - f
    if list(T1.domain.factors[df:dl + 1]) != list(T2.codomain.factors):
        raise ValueError("partial_compose: incompatible legs")
    R = tdot(T1, T2, [n1 - 1 - (df + j) for j in range(num)], list(range(num)))          # [T1 open legs, reversed(T2.domain)]

    def pos1(i):
        return i if i < f else i - num
    m = n1 - num
    cod = [pos1(i) for i in range(c1)]
    dom = ([pos1(n1 - 1 - j) for j in range(df)] + [m + (d2 - 1 - j) for j in range(d2)]
           + [pos1(n1 - 1 - j) for j in range(dl + 1, d1)])
    return permute_legs(R, cod, dom)


def tdot(A: SymmetricTensor, B: SymmetricTensor, legs1, legs2):
    """Like ``cyten.tdot`` (ops_algebra.cpp:1908-1917)."""
    l1 = A.get_leg_idcs(legs1 if isinstance(legs1, (list, tuple)) else [legs1])
    l2 = B.get_leg_idcs(legs2 if isinstance(legs2, (list, tuple)) else [legs2])
    A2 = permute_legs(A, domain=l1)
    B2 = permute_legs(B, codomain=l2)
    return compose(A2, B2)


def _decomposition_prepare(T: SymmetricTensor, new_leg_dual: bool):
    new_leg = ElementarySpace.from_largest_common_subspace([T.codomain, T.domain], is_dual=new_leg_dual)
    new_co_domain = TensorProduct([new_leg])
    combine_codomain, combine_domain = T.num_codomain_legs > 1, T.num_domain_legs > 1
    groups = []
    if combine_codomain:
        groups.append(list(range(T.num_codomain_legs)))
    if combine_domain:
        groups.append(list(range(T.num_codomain_legs, T.num_legs)))
    tens = combine_legs(T, *groups) if groups else T
    return tens, new_co_domain, combine_codomain, combine_domain


def svd(T: SymmetricTensor, new_labels=("a", "b", "c", "d"), new_leg_dual=False):
    """``T = U @ S @ Vh`` (decompositions.cpp:554-617)."""
    tens, new_co_domain, cc, cd = _decomposition_prepare(T, new_leg_dual)
    be = T.backend
    u, s, vh = be.svd(tens, new_co_domain)
    a, b, c, d = new_labels
    U = SymmetricTensor(u, tens.codomain, new_co_domain, be, [*tens.labels[:1], a])
    S = DiagonalTensor(s, new_co_domain.factors[0], be, [b, c])
    Vh = SymmetricTensor(vh, new_co_domain, tens.domain, be, [d, *tens.labels[1:]])
    if cc:
        U = split_legs(U, [0])
    if cd:
        Vh = split_legs(Vh, [1])
    return U, S, Vh


def truncated_svd(T: SymmetricTensor, new_labels=("a", "b", "c", "d"), new_leg_dual=False,
                  normalize_to=None, chi_max=None, chi_min=1, degeneracy_tol=0.0, trunc_cut=0.0,
                  svd_min=0.0):
    """decompositions.cpp:672-709: svd, truncation of the NORMALISED singular values with one global
    cut, projection of U, S, Vh.  Returns (U, S, Vh, err, renormalize)."""
    tens, new_co_domain, cc, cd = _decomposition_prepare(T, new_leg_dual)
    be = T.backend
    u, s, vh = be.svd(tens, new_co_domain)
    leg = new_co_domain.factors[0]
    S_norm = be.norm(s)
    keep, small_leg, err, new_norm = be.truncate_singular_values(
        s, leg, chi_max=chi_max, chi_min=chi_min, degeneracy_tol=degeneracy_tol, trunc_cut=trunc_cut,
        svd_min=svd_min, scale=(1.0 / S_norm if S_norm > 0 else 1.0))
    u = be.apply_keep_mask(u, 1, 1, leg, keep, small_leg)
    s = be.apply_keep_mask(s, 0, 0, leg, keep, small_leg)
    vh = be.apply_keep_mask(vh, 0, 0, leg, keep, small_leg)
    small = TensorProduct([small_leg])
    a, b, c, d = new_labels
    U = SymmetricTensor(u, tens.codomain, small, be, [*tens.labels[:1], a])
    S = DiagonalTensor(s, small_leg, be, [b, c])
    Vh = SymmetricTensor(vh, small, tens.domain, be, [d, *tens.labels[1:]])
    renormalize = 1.0
    if normalize_to is not None:
        renormalize = normalize_to / S_norm / new_norm
        be.scale_inplace(s, renormalize)
    if cc:
        U = split_legs(U, [0])
    if cd:
        Vh = split_legs(Vh, [1])
    return U, S, Vh, err, renormalize


def qr(T: SymmetricTensor, new_labels=("a", "b"), new_leg_dual=False):
    """``T = Q @ R`` (decompositions.cpp:466-508)."""
    tens, new_co_domain, cc, cd = _decomposition_prepare(T, new_leg_dual)
    be = T.backend
    q, r = be.qr(tens, new_co_domain)
    Q = SymmetricTensor(q, tens.codomain, new_co_domain, be, [*tens.labels[:1], new_labels[0]])
    R = SymmetricTensor(r, new_co_domain, tens.domain, be, [new_labels[1], *tens.labels[1:]])
    if cc:
        Q = split_legs(Q, [0])
    if cd:
        R = split_legs(R, [1])
    return Q, R


def lq(T: SymmetricTensor, new_labels=("a", "b"), new_leg_dual=False):
    """``T = L Q`` with ``Q`` an isometry from the right (``Q Q^dagger = 1``); decompositions.cpp / abelian.cpp:2304-2382.
    Computed as the QR decomposition of ``T^dagger`` (``T^dagger = Q^dagger L^dagger``): the two daggers are block
    copies, the factorisation is the batched blocked Householder kernel."""
    Qd, Rd = qr(dagger(T), new_labels=(_dual_label(new_labels[1]) if new_labels[1] else None,
                                       _dual_label(new_labels[0]) if new_labels[0] else None), new_leg_dual=new_leg_dual)
    return dagger(Rd), dagger(Qd)


def eigh(T: SymmetricTensor, new_labels=("a", "b", "c"), new_leg_dual=False, sort=None):
    """``T = V @ W @ V^dagger`` for a Hermitian T with codomain == domain (decompositions.cpp:314-424)."""
    if T.codomain != T.domain:
        raise ValueError("eigh: codomain and domain must agree")
    be = T.backend
    combine = T.num_codomain_legs > 1
    tens = T
    if combine:
        n = T.num_codomain_legs
        tens = combine_legs(T, list(range(n)), list(range(n, 2 * n)),
                            pipe_dualities=[new_leg_dual, not new_leg_dual])
    w, v, new_leg = be.eigh(tens, new_leg_dual, sort)
    W = DiagonalTensor(w, new_leg, be, [new_labels[1], new_labels[2]])
    V = SymmetricTensor(v, tens.codomain, TensorProduct([new_leg]), be, [*tens.labels[:1], new_labels[0]])
    if combine:
        V = split_legs(V, [0])
    return W, V


def norm(T) -> float:
    """Frobenius norm (src/tensors/ops_algebra.cpp:2162-2165 -> abelian.cpp:2780-2791)."""
    return T.backend.norm(T.data)


# ---- (f1) Krylov vector operations (ops_algebra.cpp:544-600, 700-760, 949-1010, 1575-1640) --------------------
def _dual_label(label):
    if label is None:
        return None
    return label[:-1] if label.endswith("*") else label + "*"


def _check_same_legs(A, B, what: str, transposed: bool = False):
    la, lb = A.legs, B.legs
    if transposed:
        lb = [l.dual for l in reversed(lb)]
    if len(la) != len(lb) or any(x != y for x, y in zip(la, lb)):
        raise ValueError(f"{what}: incompatible legs")
    if (A.num_codomain_legs != (B.num_domain_legs if transposed else B.num_codomain_legs)):
        raise ValueError(f"{what}: incompatible codomain / domain")


def inner(A: SymmetricTensor, B: SymmetricTensor, do_dagger: bool = True):
    """``<A|B>`` = trace(A^dagger B) if ``do_dagger`` else trace(A B) (B's legs = A's reversed and dualised)."""
    _check_same_legs(A, B, "inner", transposed=not do_dagger)
    return A.backend.inner(A.data, B.data, do_dagger)


def linear_combination(a, v: SymmetricTensor, b, w: SymmetricTensor) -> SymmetricTensor:
    """``a v + b w`` (tensors of equal legs; labels of ``v``)."""
    _check_same_legs(v, w, "linear_combination")
    return SymmetricTensor(v.backend.linear_combination(a, v.data, b, w.data), v.codomain, v.domain, v.backend, v.labels)


def scalar_multiply(a, T: SymmetricTensor) -> SymmetricTensor:
    return SymmetricTensor(T.backend.mul(a, T.data), T.codomain, T.domain, T.backend, T.labels)


def dagger(T):
    """Hermitian conjugate: codomain and domain swap, labels reversed and dualised (``p`` <-> ``p*``)."""
    if isinstance(T, DiagonalTensor):
        data = T.backend.lincomb_conj(T.data) if T.data.dtype == "complex128" else T.data
        return DiagonalTensor(data, T.leg, T.backend, [_dual_label(l) for l in reversed(T.labels)])
    return SymmetricTensor(T.backend.dagger(T.data), T.domain, T.codomain, T.backend,
                           [_dual_label(l) for l in reversed(T.labels)])


def scale_axis(T: SymmetricTensor, D: DiagonalTensor, leg) -> SymmetricTensor:
    """Contracts the diagonal tensor ``D`` with one leg of ``T`` without moving it (``T.legs[leg]`` must be
    ``D.leg`` or its dual): the U*S / S*Vh products of the sweep."""
    i = T.get_leg_idcs(leg)[0]
    tl = T.get_leg_co_domain(i)
    if not (tl == D.leg or tl == D.leg.dual):     # block indices count the defining sectors: same for a leg and its dual
        raise ValueError("scale_axis: incompatible legs")
    return SymmetricTensor(T.backend.scale_axis(T.data, D.data, i), T.codomain, T.domain, T.backend, T.labels)


def pinv(D: DiagonalTensor, cutoff: float = 1e-15) -> DiagonalTensor:
    """Pseudo-inverse of a diagonal tensor (``cyten.pinv``, src/tensors/ops_algebra.cpp:1505-1520)."""
    return DiagonalTensor(D.backend.cutoff_inverse(D.data, cutoff), D.leg, D.backend, D.labels)


def _install_operators():
    ST = SymmetricTensor
    ST.__add__ = lambda self, other: linear_combination(1.0, self, 1.0, other)
    ST.__sub__ = lambda self, other: linear_combination(1.0, self, -1.0, other)
    ST.__mul__ = lambda self, a: scalar_multiply(a, self)
    ST.__rmul__ = lambda self, a: scalar_multiply(a, self)
    ST.__truediv__ = lambda self, a: scalar_multiply(1.0 / a, self)
    ST.__neg__ = lambda self: scalar_multiply(-1.0, self)
    ST.hc = property(dagger)


_install_operators()

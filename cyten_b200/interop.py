"""Conversion between the reference's tensors and the B200 arena tensors.

The reference's interchange payload is ``AbelianBackendData(dtype, device, blocks, block_inds, is_sorted)``
(pybind/backends/py_abelian.cpp:95-110) on legs given by their defining sectors and
multiplicities.  These helpers are duck-typed on purpose: they never import the reference (it is not a
dependency of the product path); the caller passes the ``cyten`` module or a tensor made by it.
INTEGRATION.md shows where a cyten maintainer would call them from a ``TensorBackend`` plugin.
"""
from __future__ import annotations

import numpy as np

from .spaces import AbelianLegPipe, AbelianSymmetry, ElementarySpace, TensorProduct
from .tensors import DiagonalTensor, SymmetricTensor


def leg_from_reference(leg, symmetry: AbelianSymmetry):
    """``cyten.ElementarySpace`` / ``cyten.AbelianLegPipe`` -> the host mirror (same sector tables)."""
    sub = getattr(leg, "legs", None)
    if sub is not None and hasattr(leg, "combine_cstyle"):
        return AbelianLegPipe([leg_from_reference(l, symmetry) for l in sub], is_dual=bool(leg.is_dual),
                              combine_cstyle=bool(leg.combine_cstyle))
    ds = leg.defining_sectors
    ds = np.asarray(ds.to_numpy() if hasattr(ds, "to_numpy") else ds, dtype=np.int64).reshape(int(leg.num_sectors), -1)
    return ElementarySpace(symmetry, ds, np.asarray(leg.multiplicities, dtype=np.int64), bool(leg.is_dual))


def from_reference(T, backend, symmetry: AbelianSymmetry) -> SymmetricTensor:
    """Reference ``SymmetricTensor`` (AbelianBackend, any block backend that offers ``to_numpy``) -> arena
    tensor on ``backend``: one upload of all blocks."""
    cod = TensorProduct([leg_from_reference(l, symmetry) for l in T.codomain], symmetry=symmetry)
    dom = TensorProduct([leg_from_reference(l, symmetry) for l in T.domain], symmetry=symmetry)
    blocks = [np.asarray(b.to_numpy()) for b in T.data.blocks]
    nlegs = len(cod.factors) + len(dom.factors)
    bi = np.asarray(T.data.block_inds.to_numpy(), dtype=np.int64).reshape(len(blocks), nlegs)
    dtype = "complex128" if any(np.iscomplexobj(b) for b in blocks) or "complex" in str(T.dtype).lower() else "float64"
    return SymmetricTensor.from_numpy_blocks(blocks, bi, cod, dom, backend, labels=list(T.labels), dtype=dtype)


def diagonal_from_reference(D, backend, symmetry: AbelianSymmetry) -> DiagonalTensor:
    """Reference ``DiagonalTensor`` (blocks = vectors, block_inds rows ``[i, i]``) -> arena DiagonalTensor."""
    import torch
    from .data import AbelianBackendData
    blocks = [np.asarray(b.to_numpy()).reshape(-1) for b in D.data.blocks]
    bi = np.asarray(D.data.block_inds.to_numpy(), dtype=np.int64).reshape(len(blocks), 2)
    dtype = "complex128" if any(np.iscomplexobj(b) for b in blocks) else "float64"
    data = AbelianBackendData.empty(dtype, backend.device, bi, np.asarray([[b.size] for b in blocks], np.int64).reshape(-1, 1))
    host = np.zeros(data.buf.numel(), dtype=np.complex128 if dtype == "complex128" else np.float64)
    for b, off in zip(blocks, data.offsets):
        host[off:off + b.size] = b
    data.buf.copy_(torch.from_numpy(host))
    return DiagonalTensor(data, leg_from_reference(D.leg, symmetry), backend, list(D.labels))


def leg_to_reference(leg, ct, symmetry):
    """Host mirror -> ``ct.ElementarySpace`` / ``ct.AbelianLegPipe`` (``ct`` = the cyten module)."""
    if isinstance(leg, AbelianLegPipe):
        return ct.AbelianLegPipe([leg_to_reference(l, ct, symmetry) for l in leg.legs], is_dual=leg.is_dual,
                                 combine_cstyle=leg.combine_cstyle)
    return ct.ElementarySpace.from_defining_sectors(symmetry, leg.defining_sectors, leg.multiplicities,
                                                    is_dual=leg.is_dual)


def to_reference(T: SymmetricTensor, ct, symmetry, ref_backend=None):
    """Arena tensor -> reference ``SymmetricTensor`` on ``ref_backend`` (default abelian + numpy): one download."""
    be = ref_backend if ref_backend is not None else ct.get_backend("abelian", "numpy")
    blocks = T.data.to_numpy_blocks()
    rdt = ct.Dtype.complex128 if T.data.dtype == "complex128" else ct.Dtype.float64
    data = ct.backends.AbelianBackendData(dtype=rdt, device="cpu",
                                          blocks=[be.block_backend.as_block(np.ascontiguousarray(b)) for b in blocks],
                                          block_inds=ct.backends.BlockInds(np.ascontiguousarray(T.data.block_inds)),
                                          is_sorted=True)
    return ct.SymmetricTensor(data, [leg_to_reference(l, ct, symmetry) for l in T.codomain.factors],
                              [leg_to_reference(l, ct, symmetry) for l in T.domain.factors], backend=be,
                              labels=list(T.labels))


def diagonal_to_numpy(D) -> np.ndarray:
    """Dense diagonal (internal basis order) of a reference or arena DiagonalTensor."""
    if isinstance(D, DiagonalTensor):
        return D.to_numpy()
    out = np.zeros(int(D.leg.dim), dtype=np.float64)
    sl = np.asarray(D.leg.slices)
    bi = np.asarray(D.data.block_inds.to_numpy()).reshape(len(D.data.blocks), -1)
    for row, b in zip(bi, D.data.blocks):
        out[sl[row[0], 0]:sl[row[0], 1]] = np.real(np.asarray(b.to_numpy()))
    return out

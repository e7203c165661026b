"""Fixture loading: golden .npz (made from the real reference) -> cyten_b200 objects."""
from __future__ import annotations

from pathlib import Path

import numpy as np

from cyten_b200.spaces import AbelianLegPipe, AbelianSymmetry, ElementarySpace, TensorProduct
from cyten_b200.tensors import DiagonalTensor, SymmetricTensor

GOLDEN = Path(__file__).resolve().parent / "golden"
CASES = sorted(p.name for p in GOLDEN.glob("case_*.npz"))


def load(name):
    z = np.load(GOLDEN / name)
    return {k: z[k] for k in z.files}


def symmetry_of(fx) -> AbelianSymmetry:
    return AbelianSymmetry(tuple(int(m) for m in fx["moduli"]))


def leg_from(fx, pre, sym):
    if int(fx[pre + "kind"]) == 1:
        legs = [leg_from(fx, f"{pre}leg{i}/", sym) for i in range(int(fx[pre + "num_legs"]))]
        pipe = AbelianLegPipe(legs, is_dual=bool(fx[pre + "is_dual"]), combine_cstyle=bool(fx[pre + "combine_cstyle"]))
        # the host mirror must reproduce the reference's tables exactly
        for name in ("block_ind_map", "fusion_outcomes_sort", "block_ind_map_slices", "sector_strides"):
            assert np.array_equal(getattr(pipe, name), fx[pre + name]), (pre, name)
        assert np.array_equal(pipe.defining_sectors, fx[pre + "defining_sectors"])
        assert np.array_equal(pipe.multiplicities, fx[pre + "multiplicities"])
        return pipe
    return ElementarySpace(sym, fx[pre + "defining_sectors"], fx[pre + "multiplicities"], bool(fx[pre + "is_dual"]))


def spaces_from(fx, pre, sym):
    cod = [leg_from(fx, f"{pre}cod{i}/", sym) for i in range(int(fx[pre + "ncod"]))]
    dom = [leg_from(fx, f"{pre}dom{i}/", sym) for i in range(int(fx[pre + "ndom"]))]
    return TensorProduct(cod, symmetry=sym), TensorProduct(dom, symmetry=sym)


def blocks_from(fx, pre):
    return [fx[f"{pre}block{n}"] for n in range(int(fx[pre + "nblocks"]))]


def tensor_from(fx, pre, backend, labels=None) -> SymmetricTensor:
    sym = symmetry_of(fx)
    cod, dom = spaces_from(fx, pre, sym)
    return SymmetricTensor.from_numpy_blocks(blocks_from(fx, pre), fx[pre + "block_inds"], cod, dom, backend,
                                             labels=labels, dtype=str(fx["dtype"]))


def dense_from(fx, pre) -> np.ndarray:
    """Dense array (internal basis) of a fixture tensor."""
    sym = symmetry_of(fx)
    cod, dom = spaces_from(fx, pre, sym)
    legs = [*cod.factors, *reversed(dom.factors)]
    blocks = blocks_from(fx, pre)
    dt = np.result_type(*[b.dtype for b in blocks]) if blocks else np.float64
    out = np.zeros([l.dim for l in legs], dtype=dt)
    for row, blk in zip(fx[pre + "block_inds"], blocks):
        out[tuple(slice(*l.slices[b]) for l, b in zip(legs, row))] = blk
    return out


def assert_same_spaces(T, fx, pre):
    sym = symmetry_of(fx)
    cod, dom = spaces_from(fx, pre, sym)
    assert T.codomain == cod, pre
    assert T.domain == dom, pre


def assert_blocks_equal(T, fx, pre, exact=True, rtol=1e-12):
    """block_inds must match bit-exactly; block contents exactly (pure data movement) or to rtol."""
    assert np.array_equal(T.data.block_inds, fx[pre + "block_inds"].reshape(T.data.block_inds.shape)), pre
    got, want = T.data.to_numpy_blocks(), blocks_from(fx, pre)
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g.shape == w.shape, (pre, g.shape, w.shape)
        if exact:
            assert np.array_equal(g, w), pre
        else:
            scale = max(1.0, float(np.max(np.abs(w))) if w.size else 1.0)
            assert np.allclose(g, w, rtol=0, atol=rtol * scale * max(1, w.shape[-1] if w.ndim else 1)), pre

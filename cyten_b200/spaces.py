"""Host-side description of abelian legs: symmetry, ElementarySpace, TensorProduct, AbelianLegPipe.

These are the small integer tables the device kernels are driven by.  Names, fields and the exact
ordering conventions follow the reference (paths relative to tenpy/cyten):

* ``ElementarySpace``: ``defining_sectors`` lexsorted (last column primary), ``sector_decomposition``
  = defining sectors (or their duals for ``is_dual``), ``sector_order`` in {'sorted','dual_sorted'},
  ``multiplicities``, ``slices`` (src/symmetries/spaces.cpp:316-400, 1037-1053).
* ``TensorProduct``: sorted ``sector_decomposition`` of the fused factors + multiplicities.
* ``AbelianLegPipe``: ``sector_strides``, ``fusion_outcomes_sort``, ``block_ind_map``,
  ``block_ind_map_slices`` exactly as ``AbelianLegPipe::prepare`` builds them
  (src/symmetries/spaces.cpp:2820-2946; fields include/cyten/symmetries/spaces.h:487-494).

Only abelian group symmetries (products of U(1) and Z_N) exist here: a sector is an int vector, fusion
is component-wise addition (mod N) -- ``Symmetry::multiple_fusion_broadcast``
(src/symmetries/symmetry.cpp:375-402; U(1) add: src/symmetries/factors/u1.cpp:51).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

import numpy as np

__all__ = ["AbelianSymmetry", "u1_symmetry", "zn_symmetry", "ElementarySpace", "TensorProduct",
           "AbelianLegPipe", "make_stride", "lexsort_rows", "find_row_differences"]


def lexsort_rows(a: np.ndarray) -> np.ndarray:
    """``np.lexsort(a.T)``: stable, LAST column primary (BlockInds::lexsort_indices,
    src/backends/block_inds.cpp:290-299)."""
    a = np.asarray(a)
    if a.shape[0] == 0:
        return np.zeros(0, dtype=np.int64)
    if a.ndim == 1 or a.shape[1] == 0:
        return np.arange(a.shape[0], dtype=np.int64)
    return np.lexsort(a.T).astype(np.int64)


def find_row_differences(a: np.ndarray, include_len: bool = True) -> np.ndarray:
    """Start index of every run of equal consecutive rows (block_inds.cpp:308-336)."""
    a = np.asarray(a)
    n = a.shape[0]
    if n == 0:
        return np.zeros(1 if include_len else 0, dtype=np.int64)
    if a.ndim == 1:
        a = a[:, None]
    diff = np.ones(n, dtype=bool)
    diff[1:] = np.any(a[1:] != a[:-1], axis=1)
    res = np.nonzero(diff)[0].astype(np.int64)
    if include_len:
        res = np.concatenate([res, [n]]).astype(np.int64)
    return res


def make_stride(shape: Sequence[int], cstyle: bool = True) -> np.ndarray:
    """Strides for packing multi-indices (cyten/tools/misc.py make_stride)."""
    shape = [int(s) for s in shape]
    L = len(shape)
    stride = 1
    res = np.ones(L, dtype=np.int64)
    if cstyle:
        for a in range(L - 1, 0, -1):
            stride *= shape[a]
            res[a - 1] = stride
    else:
        for a in range(0, L - 1):
            stride *= shape[a]
            res[a + 1] = stride
    return res


@dataclass(frozen=True)
class AbelianSymmetry:
    """Product of U(1) (modulus 0) and Z_N (modulus N) factors."""
    moduli: tuple

    @property
    def sector_ind_len(self) -> int:
        return len(self.moduli)

    @property
    def trivial_sector(self) -> np.ndarray:
        return np.zeros(len(self.moduli), dtype=np.int64)

    def _reduce(self, s: np.ndarray) -> np.ndarray:
        s = np.array(s, dtype=np.int64, copy=True)
        for k, m in enumerate(self.moduli):
            if m > 0:
                s[..., k] %= m
        return s

    def dual_sectors(self, sectors: np.ndarray) -> np.ndarray:
        return self._reduce(-np.asarray(sectors, dtype=np.int64))

    def multiple_fusion_broadcast(self, parts: Sequence[np.ndarray]) -> np.ndarray:
        acc = np.array(parts[0], dtype=np.int64, copy=True)
        for p in parts[1:]:
            acc = acc + np.asarray(p, dtype=np.int64)
        return self._reduce(acc)

    def __mul__(self, other: "AbelianSymmetry") -> "AbelianSymmetry":
        return AbelianSymmetry(self.moduli + other.moduli)

    def moduli_array(self) -> np.ndarray:
        return np.asarray(self.moduli, dtype=np.int64)


u1_symmetry = AbelianSymmetry((0,))


def zn_symmetry(n: int) -> AbelianSymmetry:
    return AbelianSymmetry((int(n),))


def _as_sectors(symmetry: AbelianSymmetry, sectors) -> np.ndarray:
    s = np.asarray(sectors, dtype=np.int64)
    if s.ndim == 1:
        s = s.reshape(-1, symmetry.sector_ind_len)
    if s.shape[1] != symmetry.sector_ind_len:
        raise ValueError("sectors have the wrong number of components for this symmetry")
    return s


class ElementarySpace:
    """A leg: sectors with multiplicities (no basis_perm: the device works in the internal basis)."""

    def __init__(self, symmetry: AbelianSymmetry, defining_sectors, multiplicities=None,
                 is_dual: bool = False):
        self.symmetry = symmetry
        ds = _as_sectors(symmetry, defining_sectors)
        if multiplicities is None:
            multiplicities = np.ones(len(ds), dtype=np.int64)
        mults = np.asarray(multiplicities, dtype=np.int64)
        if mults.shape != (len(ds),):
            raise ValueError("multiplicities must have one entry per sector")
        if len(ds) > 1:
            order = lexsort_rows(ds)
            if np.any(order != np.arange(len(ds))) or len(find_row_differences(ds)) != len(ds) + 1:
                raise ValueError("defining_sectors must be sorted and duplicate-free; "
                                 "use ElementarySpace.from_defining_sectors")
        self.defining_sectors = ds
        self.multiplicities = mults
        self.is_dual = bool(is_dual)
        self.sector_decomposition = symmetry.dual_sectors(ds) if is_dual else ds
        self.sector_order = "dual_sorted" if is_dual else "sorted"
        bounds = np.concatenate([[0], np.cumsum(mults)]).astype(np.int64)
        self.slices = np.stack([bounds[:-1], bounds[1:]], axis=1)
        self.dim = int(bounds[-1])
        self.num_sectors = len(ds)

    # -- constructors mirroring the reference ---------------------------------------------------
    @classmethod
    def from_defining_sectors(cls, symmetry, defining_sectors, multiplicities=None, is_dual=False):
        """Sorts and merges duplicate sectors (reference: ElementarySpace::from_defining_sectors)."""
        ds = _as_sectors(symmetry, defining_sectors)
        mults = (np.ones(len(ds), dtype=np.int64) if multiplicities is None
                 else np.asarray(multiplicities, dtype=np.int64))
        order = lexsort_rows(ds)
        ds, mults = ds[order], mults[order]
        diffs = find_row_differences(ds)
        merged = np.add.reduceat(mults, diffs[:-1]) if len(ds) else mults
        return cls(symmetry, ds[diffs[:-1]], merged, is_dual)

    @classmethod
    def from_sector_decomposition(cls, symmetry, sector_decomposition, multiplicities=None,
                                  is_dual=False):
        sd = _as_sectors(symmetry, sector_decomposition)
        ds = symmetry.dual_sectors(sd) if is_dual else sd
        return cls.from_defining_sectors(symmetry, ds, multiplicities, is_dual)

    @classmethod
    def from_largest_common_subspace(cls, spaces, is_dual=False):
        """src/symmetries/spaces.cpp:1165-1215 (sector-wise minimum multiplicity)."""
        sp1 = spaces[0]
        if len(spaces) == 1:
            return sp1.as_ElementarySpace(is_dual)
        if len(spaces) > 2:
            pair = cls.from_largest_common_subspace(spaces[:2])
            return cls.from_largest_common_subspace([pair, *spaces[2:]], is_dual)
        sp2 = spaces[1]
        secs, mults = [], []
        lut = {tuple(s): j for j, s in enumerate(sp2.sector_decomposition.tolist())}
        for i, s in enumerate(sp1.sector_decomposition.tolist()):
            j = lut.get(tuple(s))
            if j is None:
                continue
            secs.append(s)
            mults.append(min(int(sp1.multiplicities[i]), int(sp2.multiplicities[j])))
        secs = np.asarray(secs, dtype=np.int64).reshape(-1, sp1.symmetry.sector_ind_len)
        return cls.from_sector_decomposition(sp1.symmetry, secs, mults, is_dual)

    # -- queries -----------------------------------------------------------------------------------
    @property
    def dual(self) -> "ElementarySpace":
        d = self.__dict__.get("_dual")
        if d is None:
            d = ElementarySpace(self.symmetry, self.defining_sectors, self.multiplicities, not self.is_dual)
            d.__dict__["_dual"] = self
            self.__dict__["_dual"] = d
        return d

    def as_ElementarySpace(self, is_dual: bool = False) -> "ElementarySpace":
        return ElementarySpace.from_sector_decomposition(self.symmetry, self.sector_decomposition,
                                                         self.multiplicities, is_dual)

    @property
    def sig(self) -> bytes:
        """Hashable signature (symmetry, duality, sectors, multiplicities): key material of cached plans."""
        sg = self.__dict__.get("_sig")
        if sg is None:
            import hashlib
            h = hashlib.blake2b(digest_size=16)
            h.update(repr(self.symmetry.moduli_array().tolist()).encode())
            h.update(b"d" if self.is_dual else b"n")
            h.update(self.defining_sectors.tobytes())
            h.update(self.multiplicities.tobytes())
            sg = self.__dict__["_sig"] = h.digest()
        return sg

    def sector_decomposition_where(self, sector) -> int | None:
        sector = np.asarray(sector, dtype=np.int64).reshape(-1)
        hit = np.nonzero(np.all(self.sector_decomposition == sector[None, :], axis=1))[0]
        return int(hit[0]) if len(hit) else None

    def __eq__(self, other):
        return (isinstance(other, ElementarySpace) and self.symmetry == other.symmetry
                and self.is_dual == other.is_dual
                and np.array_equal(self.defining_sectors, other.defining_sectors)
                and np.array_equal(self.multiplicities, other.multiplicities))

    __hash__ = None

    def __repr__(self):
        return (f"ElementarySpace(num_sectors={self.num_sectors}, dim={self.dim}, "
                f"is_dual={self.is_dual})")


class TensorProduct:
    """Ordered product of legs; ``sector_decomposition`` is always 'sorted'."""

    def __init__(self, factors: Sequence[ElementarySpace], symmetry: AbelianSymmetry | None = None):
        self.factors = list(factors)
        if symmetry is None:
            if not self.factors:
                raise ValueError("need a symmetry for an empty TensorProduct")
            symmetry = self.factors[0].symmetry
        self.symmetry = symmetry
        self.num_factors = len(self.factors)
        self.sector_order = "sorted"
        self._fused = None

    def _fuse(self):
        if self._fused is None:
            sym = self.symmetry
            if not self.factors:
                self._fused = (sym.trivial_sector[None, :].copy(), np.ones(1, dtype=np.int64))
            else:
                secs = self.factors[0].sector_decomposition
                mults = self.factors[0].multiplicities
                for f in self.factors[1:]:
                    s2 = sym.multiple_fusion_broadcast([np.repeat(secs, f.num_sectors, axis=0),
                                                        np.tile(f.sector_decomposition, (len(secs), 1))])
                    m2 = np.repeat(mults, f.num_sectors) * np.tile(f.multiplicities, len(mults))
                    order = lexsort_rows(s2)
                    s2, m2 = s2[order], m2[order]
                    d = find_row_differences(s2)
                    secs, mults = s2[d[:-1]], np.add.reduceat(m2, d[:-1])
                if len(self.factors) == 1:
                    order = lexsort_rows(secs)
                    secs, mults = secs[order], mults[order]
                self._fused = (secs, mults.astype(np.int64))
        return self._fused

    @property
    def sector_decomposition(self) -> np.ndarray:
        return self._fuse()[0]

    @property
    def multiplicities(self) -> np.ndarray:
        return self._fuse()[1]

    @property
    def num_sectors(self) -> int:
        return len(self._fuse()[0])

    def as_ElementarySpace(self, is_dual: bool = False) -> ElementarySpace:
        return ElementarySpace.from_sector_decomposition(self.symmetry, self.sector_decomposition,
                                                         self.multiplicities, is_dual)

    @property
    def sig(self) -> tuple:
        return tuple(f.sig for f in self.factors)

    def __getitem__(self, i):
        return self.factors[i]

    def __len__(self):
        return self.num_factors

    def __iter__(self):
        return iter(self.factors)

    def __eq__(self, other):
        return (isinstance(other, TensorProduct) and self.num_factors == other.num_factors
                and all(a == b for a, b in zip(self.factors, other.factors)))

    __hash__ = None


def _grid(strides: np.ndarray, shape: np.ndarray) -> np.ndarray:
    """grid[m, n] = (m // strides[n]) % shape[n] for m in range(prod(shape))."""
    n = int(np.prod(shape)) if len(shape) else 1
    m = np.arange(n, dtype=np.int64)[:, None]
    return (m // strides[None, :]) % shape[None, :]


class AbelianLegPipe(ElementarySpace):
    """Combined leg + the index maps that drive combine_legs / split_legs.

    Restates ``AbelianLegPipe::prepare`` (src/symmetries/spaces.cpp:2820-2946)."""

    def __init__(self, legs: Sequence[ElementarySpace], is_dual: bool = False,
                 combine_cstyle: bool = True):
        legs = list(legs)
        if not legs:
            raise ValueError("Need at least one leg")
        sym = legs[0].symmetry
        num_legs = len(legs)
        nsec = np.asarray([l.num_sectors for l in legs], dtype=np.int64)
        sector_strides = make_stride(nsec, combine_cstyle)
        grid = _grid(sector_strides, nsec)                         # (nblocks, num_legs)
        nblocks = grid.shape[0]
        mults = np.ones(nblocks, dtype=np.int64)
        parts = []
        for n, leg in enumerate(legs):
            mults *= leg.multiplicities[grid[:, n]]
            parts.append(leg.sector_decomposition[grid[:, n]])
        sectors = sym.multiple_fusion_broadcast(parts)
        if is_dual:
            sectors = sym.dual_sectors(sectors)
        sort = lexsort_rows(sectors)
        block_ind_map = np.zeros((nblocks, 3 + num_legs), dtype=np.int64)
        block_ind_map[:, 2:2 + num_legs] = grid
        block_ind_map = block_ind_map[sort]
        mults = mults[sort]
        sectors = sectors[sort]
        slices = np.concatenate([[0], np.cumsum(mults)]).astype(np.int64)
        block_ind_map[:, 0] = slices[:-1]
        block_ind_map[:, 1] = slices[1:]
        diffs = find_row_differences(sectors)
        block_starts = slices[diffs]
        unique_mults = block_starts[1:] - block_starts[:-1]
        unique_sectors = sectors[diffs[:-1]]
        J = np.repeat(np.arange(len(diffs) - 1, dtype=np.int64), np.diff(diffs))
        block_ind_map[:, 2 + num_legs] = J
        block_ind_map[:, 0] -= block_starts[J]
        block_ind_map[:, 1] -= block_starts[J]
        super().__init__(sym, unique_sectors, unique_mults, is_dual)
        self.legs = legs
        self.num_legs = num_legs
        self.combine_cstyle = bool(combine_cstyle)
        self.sector_strides = sector_strides
        self.fusion_outcomes_sort = sort.astype(np.int64)
        self.block_ind_map_slices = diffs.astype(np.int64)
        self.block_ind_map = block_ind_map
        inv = np.empty(nblocks, dtype=np.int64)
        inv[sort] = np.arange(nblocks, dtype=np.int64)
        self._inv_fusion_outcomes_sort = inv

    @property
    def dual(self) -> "AbelianLegPipe":
        """Dual legs in REVERSED order, opposite duality and combine style
        (AbelianLegPipe::dual_pipe, src/symmetries/spaces.cpp:3195-3204)."""
        d = self.__dict__.get("_dual")
        if d is None:
            d = AbelianLegPipe([l.dual for l in reversed(self.legs)], not self.is_dual, not self.combine_cstyle)
            d.__dict__["_dual"] = self
            self.__dict__["_dual"] = d
        return d

    @property
    def sig(self) -> bytes:
        sg = self.__dict__.get("_sig")
        if sg is None:
            import hashlib
            h = hashlib.blake2b(digest_size=16)
            h.update(b"pipe" + (b"d" if self.is_dual else b"n") + (b"c" if self.combine_cstyle else b"f"))
            for l in self.legs:
                h.update(l.sig)
            sg = self.__dict__["_sig"] = h.digest()
        return sg

    def __eq__(self, other):
        if isinstance(other, AbelianLegPipe):
            return (self.is_dual == other.is_dual and self.combine_cstyle == other.combine_cstyle
                    and len(self.legs) == len(other.legs)
                    and all(a == b for a, b in zip(self.legs, other.legs)))
        return ElementarySpace.__eq__(self, other)

    __hash__ = None

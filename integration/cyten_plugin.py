"""``CytenB200Backend``: the B200 kernels behind tenpy/cyten's own ``TensorBackend`` interface.

This is the drop-in boundary of SURVEY.md section 8(b), routes 1-5, with ZERO modification of cyten:

* ``cyten.backends.TensorBackend`` is bound with a trampoline (pybind/backends/py_tensor_backend.cpp:23,
  pybind/backends/py_trampolines.hpp:23-695), so this Python subclass receives every C++ virtual call that
  ``cyten.tdot`` / ``truncated_svd`` / ``qr`` / ``eigh`` / ``permute_legs`` / ``combine_legs`` / ``split_legs`` make.
* The HOT virtuals (include/cyten/backends/tensor_backend.h:122-397) run on ``cyten_b200.backend.B200AbelianBackend``:
  one device arena per tensor, plan-cached bookkeeping, the CUDA kernels of ``libcyten_b200.so``.
* The COLD virtuals (constructors, element access, masks, dense conversion ...; ~55 of them) are delegated to an inner
  reference ``AbelianBackend`` over ``ArrayApiBlockBackend(TorchNamespace)`` (src/block_backend/array_api.cpp), i.e.
  they run per block with torch on the same device.
* The interchange payload is the reference's own ``AbelianBackendData(dtype, device, blocks, block_inds, is_sorted)``
  (pybind/backends/py_abelian.cpp:95-110).  Its blocks are ``ArenaBlock`` objects (``torch_namespace.py``) that remember
  the arena they are views of, so a tensor that travels from one hot virtual to the next is recovered as the SAME arena
  without touching its bytes; blocks made by the cold path are packed into an arena once, on first hot use.
* ``install()`` seeds ``cyten._core._tensor_backend_cache[('abelian', key)]`` (src/backends/backend_factory.cpp:59-67,
  99-104), after which ``cyten.get_backend('abelian', key)`` returns the plugin.

This module imports the user's installed ``cyten``; it never imports ``oracle/``.
"""
from __future__ import annotations

import collections
import os
import sys
from pathlib import Path

import numpy as np
import torch

import cyten
from cyten.backends import AbelianBackend, TensorBackend
from cyten.backends import AbelianBackendData as RefData
from cyten.backends import BlockInds

_ROOT = Path(__file__).resolve().parents[1]
if str(_ROOT) not in sys.path:
    sys.path.insert(0, str(_ROOT))

from cyten_b200 import _lib  # noqa: E402
from cyten_b200 import tensors as b200  # noqa: E402
from cyten_b200.backend import B200AbelianBackend  # noqa: E402
from cyten_b200.data import AbelianBackendData as Arena  # noqa: E402
from cyten_b200.spaces import AbelianLegPipe, AbelianSymmetry, ElementarySpace, TensorProduct  # noqa: E402

try:
    from .torch_namespace import ArenaBlock, make_block_backend, unwrap_block
except ImportError:                                   # imported as a top-level module (pytest -p cyten_plugin)
    from torch_namespace import ArenaBlock, make_block_backend, unwrap_block

__all__ = ["CytenB200Backend", "install", "to_plugin", "to_reference"]

_TRACE = os.environ.get("CYTEN_B200_TRACE") == "1"
_STRICT = os.environ.get("CYTEN_B200_STRICT") == "1"
_HOT_DTYPES = {cyten.Dtype.float64: "float64", cyten.Dtype.complex128: "complex128"}
_REF_DTYPES = {"float64": cyten.Dtype.float64, "complex128": cyten.Dtype.complex128}


def _sector_table(sa):
    """``SectorArray`` -> numpy.  ``SectorArray.to_numpy()`` hands out the table in ~1 us; ``np.asarray(SectorArray)`` walks the
    sequence protocol element by element (~350 us for 13 sectors -- that was most of the plugin's per-call overhead)."""
    return sa.to_numpy() if hasattr(sa, "to_numpy") else np.asarray(sa)


class _NotHot(Exception):
    """The operands are outside what the kernels cover (dtype, symmetry, tensor class): use the inner backend."""


class _View:
    """What ``B200AbelianBackend`` wants from a tensor: the arena and the mirrored (co)domain."""
    __slots__ = ("data", "codomain", "domain")

    def __init__(self, data, codomain, domain):
        self.data, self.codomain, self.domain = data, codomain, domain


class CytenB200Backend(TensorBackend):
    """cyten ``TensorBackend`` whose abelian hot path runs on the B200."""

    #: virtuals served by the kernels; everything else is delegated (see ``_install_cold_virtuals``)
    HOT = ("permute_legs", "compose", "combine_legs", "split_legs", "svd", "qr", "lq", "eigh",
           "truncate_singular_values", "mask_contract_large_leg", "apply_mask_to_DiagonalTensor",
           "inner", "norm", "linear_combination", "mul", "dagger", "scale_axis", "outer", "partial_compose")

    def __init__(self, device: str = "cuda:0", _library=None):
        """``_library``: CPU test hook of ``B200AbelianBackend`` (the numpy restatement of the C-ABI); the product
        passes nothing and needs a CUDA device."""
        self.device = str(torch.device(device))
        block_backend = make_block_backend(cyten, self.device)
        super().__init__(block_backend)
        self._bb = block_backend                      # the blocks hold a raw pointer to it: keep it alive
        self.cold = AbelianBackend(block_backend)    # the reference backend serving the cold virtuals
        self.fast = B200AbelianBackend(device, _library=_library)
        self.hot_calls = collections.Counter()        # virtual name -> calls served by the kernels
        self.cold_calls = collections.Counter()       # virtual name -> calls served by the inner backend
        self.fallbacks = collections.Counter()        # HOT virtual name -> calls that fell through to the inner backend
        self.packs = 0                                # cold-path tensors packed into an arena
        self._sym_cache = {}
        self._leg_cache = {}

    # -- identification ----------------------------------------------------------------------------------------------
    def __repr__(self):
        return f"CytenB200Backend({self.device})"

    __str__ = __repr__

    def is_correct_data_type(self, data):
        return isinstance(data, RefData)

    def supports_symmetry(self, symmetry):
        return self.cold.supports_symmetry(symmetry)

    # -- symmetry / leg mirrors (integer tables only; cached by content) ------------------------------------------------
    def _sym(self, symmetry) -> AbelianSymmetry:
        key = repr(symmetry)
        sym = self._sym_cache.get(key)
        if sym is None:
            if not (symmetry.is_abelian and symmetry.has_trivial_braid):
                raise _NotHot("symmetry")
            moduli = []
            for f in symmetry.factors:
                n = getattr(f, "N", None)
                if n is not None:
                    moduli.append(int(n))
                elif "U1" in type(f).__name__ or "U(1)" in repr(f):
                    moduli.append(0)
                elif int(f.sector_ind_len) == 1 and f.num_sectors == 1:
                    moduli.append(1)                       # no symmetry: a single sector
                else:
                    raise _NotHot("symmetry factor")
            sym = self._sym_cache[key] = AbelianSymmetry(tuple(moduli))
        return sym

    def _leg(self, leg, sym):
        sub = getattr(leg, "legs", None)
        if sub is not None and hasattr(leg, "combine_cstyle"):
            subs = tuple(self._leg(l, sym) for l in sub)
            key = ("pipe", tuple(s.sig for s in subs), bool(leg.is_dual), bool(leg.combine_cstyle))
            m = self._leg_cache.get(key)
            if m is None:
                m = self._leg_cache[key] = AbelianLegPipe(list(subs), is_dual=bool(leg.is_dual),
                                                          combine_cstyle=bool(leg.combine_cstyle))
            return m
        if not hasattr(leg, "defining_sectors"):
            raise _NotHot("leg type")
        ds = np.ascontiguousarray(np.asarray(_sector_table(leg.defining_sectors), dtype=np.int64).reshape(int(leg.num_sectors), -1))
        mu = np.ascontiguousarray(np.asarray(leg.multiplicities, dtype=np.int64))
        key = (id(sym), bool(leg.is_dual), ds.tobytes(), mu.tobytes())
        m = self._leg_cache.get(key)
        if m is None:
            if len(self._leg_cache) > 8192:
                self._leg_cache.clear()
            m = self._leg_cache[key] = ElementarySpace(sym, ds, mu, bool(leg.is_dual))
        return m

    def _tp(self, tp, sym) -> TensorProduct:
        return TensorProduct([self._leg(l, sym) for l in tp.factors], symmetry=sym)

    # -- reference data <-> arena -------------------------------------------------------------------------------------
    def _arena(self, ref_data, nlegs: int) -> Arena:
        """The arena behind a reference ``AbelianBackendData``: the owner of its blocks if they are the complete,
        ordered block list of one arena (zero copy), else the blocks are packed into a new arena (once: the block
        wrappers are re-pointed at it)."""
        dtype = _HOT_DTYPES.get(ref_data.dtype)
        if dtype is None:
            raise _NotHot("dtype")
        bi = np.asarray(ref_data.block_inds.to_numpy(), dtype=np.int64)
        n = bi.shape[0] if bi.ndim == 2 else bi.size // max(nlegs, 1)
        bi = np.ascontiguousarray(bi.reshape(n, nlegs))
        if n == 0:
            return Arena.empty(dtype, self.fast.device, bi, np.zeros((0, nlegs), np.int64))
        if not _STRICT:
            # A complete, ordered arena is recognised from its first and last block (O(1); CYTEN_B200_STRICT=1 checks every
            # block).  `ref_data.blocks` converts the whole vector<Block> into Python objects (~1 us per block, several
            # hundred blocks per intermediate of the effective Hamiltonian): the two blocks are fetched by their index rows.
            first = unwrap_block(ref_data.get_block(bi[0]))
            owner = first.owner
            if owner is not None and first.index == 0 and owner.num_blocks == n and owner.dtype == dtype:
                last = first if n == 1 else unwrap_block(ref_data.get_block(bi[n - 1]))
                if (last.owner is owner and last.index == n - 1 and owner.block_inds.shape == bi.shape
                        and np.array_equal(owner.block_inds, bi)):
                    return owner
        blocks = ref_data.blocks
        first = unwrap_block(blocks[0])
        owner = first.owner
        wrapped = [unwrap_block(b) for b in blocks]
        if (owner is not None and owner.num_blocks == n and owner.dtype == dtype
                and all(w.owner is owner and w.index == i for i, w in enumerate(wrapped))):
            if owner.block_inds.shape == bi.shape and np.array_equal(owner.block_inds, bi):
                return owner
            # the same blocks under other indices (e.g. a column swap that keeps the order): alias the arena
            if not owner.is_pending and all(len(w.shape) == nlegs for w in wrapped):
                shapes = np.asarray([w.shape for w in wrapped], dtype=np.int64).reshape(n, nlegs)
                return Arena(dtype, owner.device, owner.buf, bi, owner.offsets, shapes)
        shapes = np.asarray([w.shape for w in wrapped], dtype=np.int64).reshape(n, nlegs)
        arena = Arena.empty(dtype, self.fast.device, bi, shapes)
        tdt = arena.buf.dtype
        for i, w in enumerate(wrapped):
            arena.block(i).copy_(w.t.to(tdt) if w.t.dtype != tdt else w.t)
        for i, w in enumerate(wrapped):                   # later hot calls find the arena without copying
            w.rebind(arena, i)
        self.packs += 1
        return arena

    def _ref(self, arena: Arena):
        """Arena -> reference ``AbelianBackendData`` whose blocks are lazy views into it (no bytes move)."""
        as_block = self._bb.as_block
        dev = self.device
        tdt = torch.complex128 if arena.dtype == "complex128" else torch.float64
        blocks = [as_block(ArenaBlock(None, arena, i, dev, tdt)) for i in range(arena.num_blocks)]
        bi = arena.block_inds
        if bi.ndim != 2:
            bi = bi.reshape(arena.num_blocks, -1)
        return RefData(dtype=_REF_DTYPES[arena.dtype], device=self.device, blocks=blocks,
                       block_inds=BlockInds(np.ascontiguousarray(bi)), is_sorted=True)

    def _view(self, t) -> _View:
        if not isinstance(t, cyten.SymmetricTensor) or isinstance(t, (cyten.DiagonalTensor, cyten.Mask)):
            raise _NotHot(type(t).__name__)
        sym = self._sym(t.symmetry)
        return _View(self._arena(t.data, t.num_legs), self._tp(t.codomain, sym), self._tp(t.domain, sym))

    def _diag_view(self, d):
        """DiagonalTensor -> (arena with 1-D blocks and block_inds rows [i, i], mirrored leg)."""
        if not isinstance(d, cyten.DiagonalTensor):
            raise _NotHot(type(d).__name__)
        sym = self._sym(d.symmetry)
        ref = d.data
        dtype = _HOT_DTYPES.get(ref.dtype)
        if dtype is None:
            raise _NotHot("dtype")
        blocks = ref.blocks
        n = len(blocks)
        bi = np.ascontiguousarray(np.asarray(ref.block_inds.to_numpy(), dtype=np.int64).reshape(n, 2))
        leg = self._leg(d.leg, sym)
        if n == 0:
            return Arena.empty(dtype, self.fast.device, bi, np.zeros((0, 1), np.int64)), leg
        wrapped = [unwrap_block(b) for b in blocks]
        owner = wrapped[0].owner
        if (owner is not None and owner.num_blocks == n and owner.dtype == dtype and owner.shapes.shape[1] == 1
                and all(w.owner is owner and w.index == i for i, w in enumerate(wrapped))
                and np.array_equal(owner.block_inds, bi)):
            return owner, leg
        shapes = np.asarray([w.shape for w in wrapped], dtype=np.int64).reshape(n, 1)
        arena = Arena.empty(dtype, self.fast.device, bi, shapes)
        for i, w in enumerate(wrapped):
            arena.block(i).copy_(w.t)
        for i, w in enumerate(wrapped):
            w.rebind(arena, i)
        self.packs += 1
        return arena, leg

    def _scalar(self, x):
        """python number -> ``BlockBackend::Scalar`` (a 0-d block of the carrier; what the Scalar-returning virtuals
        must hand back through the trampoline)."""
        return self._bb.as_scalar(complex(x) if isinstance(x, (complex, np.complexfloating)) else float(x))

    def _run(self, name, hot, cold):
        """Runs the kernel path; operands the kernels do not cover go to the inner backend (counted)."""
        if _TRACE:
            print(f"[cyten_b200 hot] {name}", file=sys.stderr, flush=True)
        try:
            res = hot()
        except _NotHot:
            self.fallbacks[name] += 1
            return cold()
        self.hot_calls[name] += 1
        return res

    # -- hot virtuals: leg manipulation (tensor_backend.h:125-129, 330-337, 352-355) ------------------------------------
    def permute_legs(self, a, codomain_idcs, domain_idcs, new_codomain, new_domain, mixes_codomain_domain, levels,
                     bend_right):
        def hot():
            return self._ref(self.fast.permute_legs(self._view(a), codomain_idcs, domain_idcs))
        return self._run("permute_legs", hot, lambda: self.cold.permute_legs(
            a, codomain_idcs, domain_idcs, new_codomain, new_domain, mixes_codomain_domain, levels, bend_right))

    def combine_legs(self, tensor, leg_idcs_combine, pipes, new_codomain, new_domain):
        def hot():
            v = self._view(tensor)
            sym = v.codomain.symmetry
            return self._ref(self.fast.combine_legs(v, leg_idcs_combine, [self._leg(p, sym) for p in pipes],
                                                    self._tp(new_codomain, sym), self._tp(new_domain, sym)))
        return self._run("combine_legs", hot, lambda: self.cold.combine_legs(tensor, leg_idcs_combine, pipes,
                                                                              new_codomain, new_domain))

    def split_legs(self, a, leg_idcs, new_codomain, new_domain):
        def hot():
            v = self._view(a)
            sym = v.codomain.symmetry
            return self._ref(self.fast.split_legs(v, leg_idcs, self._tp(new_codomain, sym), self._tp(new_domain, sym)))
        return self._run("split_legs", hot, lambda: self.cold.split_legs(a, leg_idcs, new_codomain, new_domain))

    # -- hot virtuals: contraction (tensor_backend.h:132, 313-320) -------------------------------------------------------
    def compose(self, a, b):
        return self._run("compose", lambda: self._ref(self.fast.compose(self._view(a), self._view(b))),
                         lambda: self.cold.compose(a, b))

    def _as_b200(self, t) -> b200.SymmetricTensor:
        v = self._view(t)
        return b200.SymmetricTensor(v.data, v.codomain, v.domain, self.fast, [None] * t.num_legs)

    def outer(self, a, b):
        return self._run("outer", lambda: self._ref(b200.outer(self._as_b200(a), self._as_b200(b)).data),
                         lambda: self.cold.outer(a, b))

    def partial_compose(self, a, b, a_first_leg, new_codomain, new_domain):
        return self._run("partial_compose",
                         lambda: self._ref(b200.partial_compose(self._as_b200(a), self._as_b200(b), int(a_first_leg)).data),
                         lambda: self.cold.partial_compose(a, b, a_first_leg, new_codomain, new_domain))

    # -- hot virtuals: factorisations (tensor_backend.h:185-188, 270-271, 340-341, 362-364) ----------------------------
    def svd(self, a, new_co_domain, algorithm):
        def hot():
            v = self._view(a)
            u, s, vh = self.fast.svd(v, self._tp(new_co_domain, v.codomain.symmetry), algorithm)
            return self._ref(u), self._ref(s), self._ref(vh)
        return self._run("svd", hot, lambda: self.cold.svd(a, new_co_domain, algorithm))

    def qr(self, a, new_co_domain):
        def hot():
            v = self._view(a)
            q, r = self.fast.qr(v, self._tp(new_co_domain, v.codomain.symmetry))
            return self._ref(q), self._ref(r)
        return self._run("qr", hot, lambda: self.cold.qr(a, new_co_domain))

    def lq(self, tensor, new_co_domain):
        def hot():
            # T = L Q  <=>  T^dagger = Q^dagger L^dagger: the QR factorisation of the dagger (abelian.cpp:2304-2382)
            v = self._view(tensor)
            vd = _View(self.fast.dagger(v.data), v.domain, v.codomain)
            qd, rd = self.fast.qr(vd, self._tp(new_co_domain, v.codomain.symmetry))
            return self._ref(self.fast.dagger(rd)), self._ref(self.fast.dagger(qd))
        return self._run("lq", hot, lambda: self.cold.lq(tensor, new_co_domain))

    def eigh(self, a, new_leg_dual, sort):
        def hot():
            v = self._view(a)
            w, vv, _ = self.fast.eigh(v, bool(new_leg_dual), sort)
            return self._ref(w), self._ref(vv), a.domain.as_ElementarySpace(bool(new_leg_dual))
        return self._run("eigh", hot, lambda: self.cold.eigh(a, new_leg_dual, sort))

    # -- hot virtuals: truncation (tensor_backend.h:122, 279-285, 390-397) -----------------------------------------------
    def truncate_singular_values(self, S, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error):
        def hot():
            s, leg = self._diag_view(S)
            keep, small, err, new_norm = self.fast.truncate_singular_values(
                s, leg, chi_max=chi_max, chi_min=chi_min, degeneracy_tol=degeneracy_tol, trunc_cut=trunc_cut,
                svd_min=svd_min, minimize_error=minimize_error)
            return (*self._mask_from_keep(keep, S.leg), float(err), float(new_norm))
        return self._run("truncate_singular_values", hot, lambda: self.cold.truncate_singular_values(
            S, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error))

    def _mask_from_keep(self, keep: np.ndarray, large_leg):
        """``mask_from_block`` (abelian.cpp:2584-2639) for a keep mask that is already on the host: bool blocks per kept
        sector (device views + a host copy riding on the block wrapper), block_inds rows [small, large], small leg."""
        keep = np.asarray(keep, dtype=bool)
        keep_dev = torch.from_numpy(keep).to(self.fast.device) if len(keep) else torch.zeros(0, dtype=torch.bool)
        slices = np.asarray(large_leg.slices).reshape(-1, 2)
        defining = np.asarray(_sector_table(large_leg.defining_sectors), dtype=np.int64).reshape(len(slices), -1)
        perm = large_leg._basis_perm
        perm = None if perm is None else np.asarray(large_leg.basis_perm)
        blocks, rows, sectors, mults, ranks = [], [], [], [], []
        for i, (b0, b1) in enumerate(slices):
            k = keep[b0:b1]
            m = int(k.sum())
            if m == 0:
                continue
            blk = ArenaBlock(keep_dev[b0:b1])
            blk.host = k
            blocks.append(self._bb.as_block(blk, cyten.Dtype.bool))
            rows.append((len(rows), i))
            sectors.append(defining[i])
            mults.append(m)
            if perm is not None:
                ranks.append(perm[b0:b1][k])
        if blocks:
            bi = np.asarray(rows, dtype=np.int64)
            secs = np.asarray(sectors, dtype=np.int64)
            basis_perm = None if perm is None else np.argsort(np.argsort(np.concatenate(ranks))).tolist()
        else:
            bi = np.zeros((0, 2), np.int64)
            secs = np.zeros((0, defining.shape[1]), np.int64)
            basis_perm = None
        data = RefData(dtype=cyten.Dtype.bool, device=self.device, blocks=blocks, block_inds=BlockInds(bi), is_sorted=True)
        small_leg = cyten.ElementarySpace(large_leg.symmetry, secs, np.asarray(mults, dtype=np.int64),
                                          bool(large_leg.is_dual), basis_perm)
        return data, small_leg

    def _keep_of_mask(self, mask) -> np.ndarray:
        """Host keep vector over the large leg of a projection Mask."""
        large = mask.large_leg
        slices = np.asarray(large.slices).reshape(-1, 2)
        keep = np.zeros(int(large.dim), dtype=bool)
        md = mask.data
        n = len(md.blocks)
        bi = np.asarray(md.block_inds.to_numpy(), dtype=np.int64).reshape(n, 2)
        col = 1 if mask.is_projection else 0
        for row, b in zip(bi, md.blocks):
            w = unwrap_block(b)
            host = getattr(w, "host", None)
            if host is None:
                host = w.t.cpu().numpy().astype(bool)
            b0, b1 = slices[row[col]]
            keep[b0:b1] = host
        return keep

    def mask_contract_large_leg(self, tensor, mask, leg_idx):
        def hot():
            v = self._view(tensor)
            n = tensor.num_legs
            li = int(leg_idx) % n
            if n != 2 and li not in (0, n - 1):
                raise _NotHot("leg in the middle")
            in_domain = li >= tensor.num_codomain_legs
            sym = v.codomain.symmetry
            ref_small = mask.small_leg
            small = self._leg(ref_small, sym)
            large = self._leg(mask.large_leg, sym)
            keep = self._keep_of_mask(mask)
            data = v.data
            shapes = data.shapes
            # the block as a matrix with the masked leg as its first or last axis
            if li == 0:
                two = Arena(data.dtype, data.device, data.buf, data.block_inds,
                            data.offsets, np.stack([shapes[:, 0], np.prod(shapes[:, 1:], axis=1)], axis=1))
                axis = 0
            else:
                two = Arena(data.dtype, data.device, data.buf, data.block_inds,
                            data.offsets, np.stack([np.prod(shapes[:, :-1], axis=1), shapes[:, -1]], axis=1))
                axis = 1
            res = self.fast.apply_keep_mask(two, axis, li, large, keep, small)
            legs = [*v.codomain.factors, *reversed(v.domain.factors)]
            legs[li] = small
            full = np.stack([np.asarray(l.multiplicities)[res.block_inds[:, j]] for j, l in enumerate(legs)], axis=1)
            res = Arena(res.dtype, res.device, res.buf, res.block_inds, res.offsets, full.reshape(res.num_blocks, n))
            factors = list(tensor.domain.factors if in_domain else tensor.codomain.factors)
            pos = (n - 1 - li) if in_domain else li
            factors[pos] = ref_small
            tp = cyten.TensorProduct(factors, tensor.symmetry)
            return (self._ref(res), tensor.codomain, tp) if in_domain else (self._ref(res), tp, tensor.domain)
        return self._run("mask_contract_large_leg", hot, lambda: self.cold.mask_contract_large_leg(tensor, mask, leg_idx))

    def apply_mask_to_DiagonalTensor(self, tensor, mask):
        def hot():
            s, leg = self._diag_view(tensor)
            sym = leg.symmetry
            small = self._leg(mask.small_leg, sym)
            keep = self._keep_of_mask(mask)
            return self._ref(self.fast.apply_keep_mask(s, 0, 0, leg, keep, small))
        return self._run("apply_mask_to_DiagonalTensor", hot, lambda: self.cold.apply_mask_to_DiagonalTensor(tensor, mask))

    # -- hot virtuals: the Krylov vector operations (tensor_backend.h:137, 251-253, 265-268, 307, 310, 349) ---------------
    def _any_arena(self, t) -> Arena:
        if isinstance(t, cyten.DiagonalTensor):
            return self._diag_view(t)[0]
        if isinstance(t, cyten.SymmetricTensor) and not isinstance(t, cyten.Mask):
            self._sym(t.symmetry)
            return self._arena(t.data, t.num_legs)
        raise _NotHot(type(t).__name__)

    def inner(self, a, b, do_dagger):
        def hot():
            x, y = self._view(a), self._view(b)
            return self._scalar(self.fast.inner(x.data, y.data, bool(do_dagger)))
        return self._run("inner", hot, lambda: self.cold.inner(a, b, do_dagger))

    def norm(self, a):
        return self._run("norm", lambda: self._scalar(float(self.fast.norm(self._any_arena(a)))), lambda: self.cold.norm(a))

    def linear_combination(self, a, v, b, w):
        def hot():
            x, y = self._view(v), self._view(w)
            return self._ref(self.fast.linear_combination(_num(a), x.data, _num(b), y.data))
        return self._run("linear_combination", hot, lambda: self.cold.linear_combination(a, v, b, w))

    def mul(self, a, b):
        def hot():
            return self._ref(self.fast.mul(_num(a), self._view(b).data))
        return self._run("mul", hot, lambda: self.cold.mul(a, b))

    def dagger(self, a):
        return self._run("dagger", lambda: self._ref(self.fast.dagger(self._view(a).data)), lambda: self.cold.dagger(a))

    def scale_axis(self, a, b, leg):
        def hot():
            x = self._view(a)
            d, _ = self._diag_view(b)
            return self._ref(self.fast.scale_axis(x.data, d, int(leg) % a.num_legs))
        return self._run("scale_axis", hot, lambda: self.cold.scale_axis(a, b, leg))


def _partial_trace(self, tensor, pairs, levels):
    """Cold virtual with one adaptation: for a full trace the reference returns ``(data, None, None)``
    (abelian.cpp:3042-3054), and ``None`` does not convert to ``TensorProduct::Ptr`` on the way back through the
    trampoline's tuple caster; the caller ignores both spaces in that case (ops_algebra.cpp:1488-1491)."""
    self.cold_calls["partial_trace"] += 1
    data, codomain, domain = self.cold.partial_trace(tensor, pairs, levels)
    if not isinstance(data, RefData):
        # the Python binding of the inner backend hands a full trace back as a Scalar; the trampoline wants the
        # single-block data of abelian.cpp:3044-3053
        val = data.to_numpy()
        tdt = torch.complex128 if np.iscomplexobj(val) else torch.float64
        blk = self._bb.as_block(ArenaBlock(torch.tensor(val.item(), dtype=tdt, device=self.fast.device)))
        data = RefData(dtype=cyten.Dtype.complex128 if tdt == torch.complex128 else cyten.Dtype.float64, device=self.device,
                       blocks=[blk], block_inds=BlockInds(np.zeros((1, 0), np.int64)), is_sorted=True)
    if codomain is None:
        codomain = cyten.TensorProduct([], tensor.symmetry)
    if domain is None:
        domain = cyten.TensorProduct([], tensor.symmetry)
    return data, codomain, domain


CytenB200Backend.partial_trace = _partial_trace


def _num(x):
    """cyten Scalar / numpy scalar / python number -> python number."""
    if hasattr(x, "to_numpy"):
        x = x.to_numpy()
    if isinstance(x, np.ndarray):
        x = x.item()
    if isinstance(x, (complex, np.complexfloating)):
        x = complex(x)
        return x.real if x.imag == 0.0 else x
    return float(x)


def _install_cold_virtuals():
    """Every other ``TensorBackend`` virtual is a counted pass-through to the inner reference ``AbelianBackend``
    (the trampoline looks the override up on the Python type, so each one must exist as a method)."""
    skip = set(CytenB200Backend.HOT) | {"block_backend", "can_decompose_tensors", "Data", "item", "is_correct_data_type",
                                        "supports_symmetry"}
    for name in dir(TensorBackend):
        if name.startswith("__") or name in skip or name in CytenB200Backend.__dict__:
            continue
        if name.startswith("_pybind11"):
            continue

        def make(name):
            def method(self, *args, **kwargs):
                self.cold_calls[name] += 1
                if _TRACE:
                    print(f"[cyten_b200 cold] {name}", file=sys.stderr, flush=True)
                return getattr(self.cold, name)(*args, **kwargs)
            method.__name__ = name
            method.__qualname__ = f"CytenB200Backend.{name}"
            method.__doc__ = f"Cold virtual: delegated to the reference AbelianBackend.{name}."
            return method
        setattr(CytenB200Backend, name, make(name))


_install_cold_virtuals()


# ---------------------------------------------------------------------------------------------------------------------
# registration and conversion
# ---------------------------------------------------------------------------------------------------------------------
_INSTANCES: dict = {}


def install(device: str = "cuda:0", key: str = "b200", _library=None) -> CytenB200Backend:
    """Makes ``cyten.get_backend('abelian', key)`` return the plugin (backend_factory.cpp:59-67, 99-104: the factory
    consults ``cyten._core._tensor_backend_cache`` first; the dict is created lazily, so create it if absent)."""
    core = cyten._core
    cache = getattr(core, "_tensor_backend_cache", None)
    if cache is None:
        cyten.get_backend("abelian", "numpy")              # first call creates the cache dict
        cache = getattr(core, "_tensor_backend_cache", None)
    if cache is None:
        cache = {}
        core._tensor_backend_cache = cache
    backend = _INSTANCES.get((device, key))
    if backend is None:
        backend = _INSTANCES[(device, key)] = CytenB200Backend(device, _library=_library)
    cache[("abelian", key)] = backend
    return backend


def to_plugin(T, backend: CytenB200Backend):
    """Reference tensor on any abelian backend -> the same tensor on the plugin (one upload of all blocks).
    ``Tensor.to_backend`` cannot do this: it dispatches on ``dynamic_pointer_cast<AbelianBackend>``
    (src/tensors/symmetric_tensor.cpp:664-720)."""
    as_block = backend._bb.as_block
    dev = backend.fast.device
    if isinstance(T, cyten.DiagonalTensor):
        blocks = [as_block(ArenaBlock(torch.from_numpy(np.ascontiguousarray(b.to_numpy())).to(dev))) for b in T.data.blocks]
        data = RefData(dtype=T.data.dtype, device=backend.device, blocks=blocks, block_inds=T.data.block_inds, is_sorted=True)
        return cyten.DiagonalTensor(data, T.leg, backend=backend, labels=list(T.labels))
    blocks = [as_block(ArenaBlock(torch.from_numpy(np.array(b.to_numpy(), order="C", copy=True)).to(dev)))
              for b in T.data.blocks]
    data = RefData(dtype=T.data.dtype, device=backend.device, blocks=blocks, block_inds=T.data.block_inds, is_sorted=True)
    return cyten.SymmetricTensor(data, list(T.codomain.factors), list(T.domain.factors), backend=backend,
                                 labels=list(T.labels))


def to_reference(T, ref_backend=None):
    """Plugin tensor -> the same tensor on the reference's abelian + numpy backend (one download)."""
    be = ref_backend if ref_backend is not None else cyten.get_backend("abelian", "numpy")
    as_block = be.block_backend.as_block
    blocks = [as_block(np.ascontiguousarray(np.asarray(unwrap_block(b)))) for b in T.data.blocks]
    data = RefData(dtype=T.data.dtype, device="cpu", blocks=blocks, block_inds=T.data.block_inds, is_sorted=True)
    if isinstance(T, cyten.DiagonalTensor):
        return cyten.DiagonalTensor(data, T.leg, backend=be, labels=list(T.labels))
    return cyten.SymmetricTensor(data, list(T.codomain.factors), list(T.domain.factors), backend=be, labels=list(T.labels))

"""Device block arena: the B200 counterpart of the reference's ``AbelianBackendData``.

Reference: ``AbelianBackendData{dtype, device, blocks, block_inds}`` with ``block_inds`` kept
``np.lexsort(block_inds.T)``-sorted (include/cyten/backends/abelian.h:51-83,
src/backends/abelian.cpp:57-81).  Here all blocks of a tensor live in ONE device allocation (the
arena, owned by a torch tensor -- PyTorch is only the allocator); block ``n`` is the C-contiguous
array of shape ``shapes[n]`` starting at element ``offsets[n]`` (16-element = 128-byte aligned for
float64).  ``block_inds`` / ``offsets`` / ``shapes`` have a host copy (numpy int64, the interchange
format) and are uploaded where a kernel needs them.
"""
from __future__ import annotations

import ctypes as C
import hashlib

import numpy as np
import torch

from . import _lib

ALIGN = 16  # elements

_TORCH_DTYPES = {"float64": torch.float64, "complex128": torch.complex128}
_NP_DTYPES = {"float64": np.float64, "complex128": np.complex128}
_CYB_DTYPES = {"float64": _lib.CYB_F64, "complex128": _lib.CYB_C128}


def dtype_name(dtype) -> str:
    if isinstance(dtype, str):
        name = dtype
    elif isinstance(dtype, torch.dtype):
        name = {torch.float64: "float64", torch.complex128: "complex128"}.get(dtype)
    else:
        name = np.dtype(dtype).name
    if name not in _TORCH_DTYPES:
        raise ValueError(f"unsupported dtype {dtype!r}: the B200 path computes in float64/complex128")
    return name


def aligned_offsets(sizes: np.ndarray) -> tuple[np.ndarray, int]:
    """Element offsets of consecutive blocks, each start aligned to ALIGN elements."""
    sizes = np.asarray(sizes, dtype=np.int64)
    padded = (sizes + ALIGN - 1) // ALIGN * ALIGN
    offs = np.zeros(len(sizes), dtype=np.int64)
    if len(sizes):
        offs[1:] = np.cumsum(padded)[:-1]
    return offs, int(padded.sum())


def ptr(t) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(device=None) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def to_device_i64(a: np.ndarray, device) -> torch.Tensor:
    """Uploads a small int64 table."""
    a = np.ascontiguousarray(a, dtype=np.int64)
    return torch.from_numpy(a).to(device, non_blocking=False)


class AbelianBackendData:
    """All blocks of one tensor in a single device arena + the integer tables describing them."""

    __slots__ = ("dtype", "device", "_buf", "block_inds", "offsets", "shapes", "_dev_tables", "_sig",
                 "_pending", "_relayout")

    def __init__(self, dtype, device, buf, block_inds, offsets, shapes, sig=None, pending=None):
        self.dtype = dtype_name(dtype)
        self.device = torch.device(device)
        self._buf = buf
        self.block_inds = np.ascontiguousarray(block_inds, dtype=np.int64)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        self.shapes = np.ascontiguousarray(shapes, dtype=np.int64)
        self._dev_tables = {}
        self._sig = sig
        # deferred layout change (permute_legs): a callable producing the arena on first use.  The
        # reference's numpy blocks are lazy views in the same place (numpy.cpp permute_axes); here the
        # consumer decides where the one copy happens (compose folds it into its operand re-layout).
        self._pending = pending
        self._relayout = None   # cache of re-laid-out copies of this (immutable) arena, see backend.compose

    @property
    def buf(self) -> torch.Tensor:
        if self._buf is None:
            buf = self._pending[0]()
            if getattr(buf, "_cyb_dummy", False):
                return buf             # plan-only (dry-run) pass: nothing was computed, the permutation stays pending
            self._buf, self._pending = buf, None
        return self._buf

    @buf.setter
    def buf(self, value):
        self._buf, self._pending, self._relayout = value, None, None

    @property
    def is_pending(self) -> bool:
        return self._buf is None

    @property
    def pending_source(self):
        """(parent data, axes_perm, row order) of an unmaterialised permutation, else None."""
        return self._pending[1] if self._buf is None else None

    @property
    def sig(self) -> bytes:
        """Structure signature (dtype + integer tables, not the contents): the key of cached plans."""
        if self._sig is None:
            h = hashlib.blake2b(digest_size=16)
            h.update(self.dtype.encode())
            h.update(np.asarray(self.block_inds.shape, dtype=np.int64).tobytes())
            h.update(self.block_inds.tobytes())
            h.update(self.shapes.tobytes())
            h.update(self.offsets.tobytes())
            self._sig = h.digest()
        return self._sig

    # -- construction ------------------------------------------------------------------------------
    @classmethod
    def empty(cls, dtype, device, block_inds, shapes, zero=False):
        block_inds = np.asarray(block_inds, dtype=np.int64)
        nlegs = block_inds.shape[1] if block_inds.ndim == 2 else 0
        shapes = np.asarray(shapes, dtype=np.int64)
        shapes = shapes.reshape(len(block_inds), -1) if len(block_inds) else shapes.reshape(0, nlegs)   # no blocks at all
        sizes = np.prod(shapes, axis=1).astype(np.int64) if shapes.shape[1] else np.ones(len(shapes), np.int64)
        offs, total = aligned_offsets(sizes)
        tdt = _TORCH_DTYPES[dtype_name(dtype)]
        buf = (torch.zeros if zero else torch.empty)(max(total, ALIGN), dtype=tdt, device=device)
        return cls(dtype, device, buf, block_inds, offs, shapes)

    @classmethod
    def from_numpy_blocks(cls, blocks, block_inds, device, dtype=None, is_sorted=False):
        """Uploads host blocks (e.g. the reference's ``data.blocks``); lexsorts unless ``is_sorted``."""
        block_inds = np.asarray(block_inds, dtype=np.int64)
        nlegs = block_inds.shape[1] if block_inds.ndim == 2 else 0
        block_inds = block_inds.reshape(len(blocks), nlegs)
        if dtype is None:
            dtype = np.result_type(*[np.asarray(b).dtype for b in blocks]) if blocks else np.float64
        name = dtype_name(dtype)
        if not is_sorted and len(blocks) > 1:
            from .spaces import lexsort_rows
            order = lexsort_rows(block_inds)
            block_inds = block_inds[order]
            blocks = [blocks[i] for i in order]
        shapes = np.asarray([np.asarray(b).shape for b in blocks], dtype=np.int64).reshape(len(blocks), nlegs)
        res = cls.empty(name, device, block_inds, shapes)
        host = np.zeros(res.buf.numel(), dtype=_NP_DTYPES[name])
        for b, off in zip(blocks, res.offsets):
            b = np.ascontiguousarray(b, dtype=_NP_DTYPES[name])
            host[off:off + b.size] = b.reshape(-1)
        res.buf.copy_(torch.from_numpy(host))
        return res

    # -- views -------------------------------------------------------------------------------------
    @property
    def num_blocks(self) -> int:
        return len(self.offsets)

    @property
    def cyb_dtype(self) -> int:
        return _CYB_DTYPES[self.dtype]

    @property
    def sizes(self) -> np.ndarray:
        if self.shapes.shape[1] == 0:
            return np.ones(len(self.shapes), dtype=np.int64)
        return np.prod(self.shapes, axis=1).astype(np.int64)

    def block(self, n: int) -> torch.Tensor:
        shape = tuple(int(s) for s in self.shapes[n])
        size = int(np.prod(shape)) if shape else 1
        off = int(self.offsets[n])
        return self.buf[off:off + size].view(shape)

    @property
    def blocks(self):
        return [self.block(n) for n in range(self.num_blocks)]

    def to_numpy_blocks(self):
        host = self.buf.cpu().numpy()
        out = []
        for n in range(self.num_blocks):
            shape = tuple(int(s) for s in self.shapes[n])
            size = int(np.prod(shape)) if shape else 1
            off = int(self.offsets[n])
            out.append(host[off:off + size].reshape(shape).copy())
        return out

    def dev(self, name: str) -> torch.Tensor:
        """Cached device copy of one of the integer tables ('block_inds', 'offsets', 'shapes')."""
        t = self._dev_tables.get(name)
        if t is None:
            t = to_device_i64(getattr(self, name), self.device)
            self._dev_tables[name] = t
        return t

    def nbytes(self) -> int:
        return int(self.sizes.sum()) * (16 if self.dtype == "complex128" else 8)

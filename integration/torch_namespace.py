"""Array-API namespace over device-resident torch tensors: the block carrier of the cyten plugin.

The reference offers exactly one way to put foreign device arrays behind its ``BlockBackend`` interface without a host
round trip: ``ArrayApiBlockBackend(api_namespace, default_device)`` (src/block_backend/array_api.cpp:405-431).  Its
``as_block(obj)`` calls ``api.asarray(obj)`` and then ``1.0 * obj`` (:569-597), and every cold-path block operation is a
call into ``api`` (``abs, matmul, reshape, permute_dims, zeros, linalg.svd ...``).

``ArenaBlock`` is the array object: a thin Python wrapper around a torch tensor that is (for results of the hot path) a
VIEW INTO THE ARENA of a ``cyten_b200.data.AbelianBackendData`` -- ``owner`` / ``index`` remember which, so that the
plugin recovers the whole arena from a reference tensor without copying (``unwrap``).  The view is created lazily: a
tensor that only travels from one hot virtual to the next never builds per-block torch views.

``asarray`` and ``1.0 * block`` return the SAME object (zero copy); everything else computes with torch on the
block's device and returns a fresh, arena-less ``ArenaBlock``.  This module is cold-path glue: none of it runs inside
the hot virtuals.
"""
from __future__ import annotations

import numpy as np
import torch

__all__ = ["ArenaBlock", "TorchNamespace", "unwrap_block", "NAMESPACE"]


def _t(x):
    """ArenaBlock -> torch tensor; everything else unchanged."""
    return x.t if isinstance(x, ArenaBlock) else x


def _from_np(a: np.ndarray) -> torch.Tensor:
    """numpy -> torch without promoting 0-d arrays to 1-d (``np.ascontiguousarray`` does) and without sharing
    read-only memory."""
    if a.ndim == 0:
        return torch.as_tensor(a.item(), dtype=_NP2T.get(a.dtype.name))
    if not (a.flags.c_contiguous and a.flags.writeable):
        a = np.array(a, order="C", copy=True)
    return torch.from_numpy(a)


_NP2T = {"float64": torch.float64, "float32": torch.float32, "complex128": torch.complex128, "complex64": torch.complex64,
         "bool": torch.bool, "int64": torch.int64, "int32": torch.int32}


def _tt(x, like=None):
    """Operand of a binary operation -> something torch accepts next to ``like``."""
    if isinstance(x, ArenaBlock):
        return x.t
    if isinstance(x, np.ndarray):
        t = _from_np(x)
        return t.to(like.device) if like is not None else t
    if isinstance(x, np.generic):
        return x.item()
    return x


_LAST_SHAPE_READ = [None]


class _Unwrap:
    """Key for ``block[key]``: the reference forwards unknown keys to ``obj.__getitem__`` (array_api.cpp:181-189), which
    deposits the wrapped object here."""
    __slots__ = ("obj",)

    def __init__(self):
        self.obj = None


class ArenaBlock:
    """One dense block.  ``owner``/``index``: the arena data object and block number this block is a view of (or
    ``None`` for blocks produced by cold-path arithmetic).  ``device`` / ``dtype`` are plain attributes (the reference
    reads them several times per block when it wraps the object, array_api.cpp:112-122)."""
    __slots__ = ("_tensor", "owner", "index", "host", "device", "dtype", "__weakref__")
    __array_priority__ = 1000

    def __init__(self, tensor=None, owner=None, index=-1, device=None, dtype=None):
        self._tensor = tensor
        self.owner = owner
        self.index = index
        self.host = None      # optional host copy of a small block (keep masks), spares a device read-back
        if tensor is not None:
            self.device = str(tensor.device)
            self.dtype = tensor.dtype
        else:
            self.device = device if device is not None else str(owner.device)
            self.dtype = dtype if dtype is not None else (torch.complex128 if owner.dtype == "complex128" else torch.float64)

    def rebind(self, owner, index):
        """Points this block at block ``index`` of the arena ``owner`` (same values, see CytenB200Backend._arena)."""
        self._tensor, self.owner, self.index = None, owner, index
        self.device = str(owner.device)

    # -- the torch view (lazy for arena blocks) -------------------------------------------------------------------
    @property
    def t(self) -> torch.Tensor:
        if self._tensor is None:
            self._tensor = self.owner.block(self.index)
        return self._tensor

    # -- array attributes the reference reads ---------------------------------------------------------------------
    @property
    def shape(self):
        _LAST_SHAPE_READ[0] = self        # see unwrap_block
        if self._tensor is None:
            return tuple(int(s) for s in self.owner.shapes[self.index])
        return tuple(self._tensor.shape)

    @property
    def ndim(self):
        return len(self.shape)

    @property
    def size(self):
        n = 1
        for s in self.shape:
            n *= s
        return n

    def item(self):
        return self.t.item()

    def __array__(self, dtype=None, copy=None):
        a = self.t.detach().cpu().numpy()
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return self.shape[0]

    def __float__(self):
        return float(self.t.item())

    def __complex__(self):
        return complex(self.t.item())

    def __int__(self):
        return int(self.t.item())

    def __index__(self):
        return int(self.t.item())

    def __bool__(self):
        return bool(self.t.item())

    def __repr__(self):
        return repr(np.asarray(self))         # like the numpy blocks of the reference (its reprs embed this)

    # -- indexing -------------------------------------------------------------------------------------------------
    @staticmethod
    def _key(key):
        if isinstance(key, tuple):
            return tuple(ArenaBlock._key1(k) for k in key)
        return ArenaBlock._key1(key)

    @staticmethod
    def _key1(k):
        if isinstance(k, ArenaBlock):
            return k.t
        if isinstance(k, np.ndarray):
            return _from_np(k)
        if isinstance(k, (list,)) and k and isinstance(k[0], (bool, np.bool_)):
            return torch.as_tensor(k, dtype=torch.bool)
        return k

    def __getitem__(self, key):
        if isinstance(key, _Unwrap):
            key.obj = self
            return self
        return ArenaBlock(self.t[self._key(key)])

    def __setitem__(self, key, value):
        v = _tt(value, self.t)
        if isinstance(v, torch.Tensor) and v.dtype != self.t.dtype:
            v = v.to(self.t.dtype)
        self.t[self._key(key)] = v

    # -- arithmetic ---------------------------------------------------------------------------------------------------
    def _bin(self, other, op, reverse=False):
        a = self.t
        b = _tt(other, a)
        if isinstance(b, torch.Tensor) and b.device != a.device:
            b = b.to(a.device)
        try:
            return ArenaBlock(op(b, a) if reverse else op(a, b))
        except TypeError:
            return NotImplemented

    def __add__(self, o):
        return self._bin(o, torch.add)

    def __radd__(self, o):
        return self._bin(o, torch.add, True)

    def __sub__(self, o):
        return self._bin(o, torch.sub)

    def __rsub__(self, o):
        return self._bin(o, torch.sub, True)

    def __mul__(self, o):
        return self._bin(o, torch.mul)

    def __rmul__(self, o):
        # ``1.0 * block`` is how ArrayApiBlockBackend::as_block forces a float dtype (array_api.cpp:594-595): for a
        # block that already is float64 / complex128 the same object comes back -- zero copy
        if o == 1.0 and type(o) is float and (self.dtype is torch.float64 or self.dtype is torch.complex128):
            return self
        return self._bin(o, torch.mul, True)

    def __truediv__(self, o):
        return self._bin(o, torch.true_divide)

    def __rtruediv__(self, o):
        return self._bin(o, torch.true_divide, True)

    def __pow__(self, o):
        return self._bin(o, torch.pow)

    def __rpow__(self, o):
        return self._bin(o, torch.pow, True)

    def __neg__(self):
        return ArenaBlock(-self.t)

    def __abs__(self):
        return ArenaBlock(torch.abs(self.t))

    def __matmul__(self, o):
        return self._bin(o, torch.matmul)

    def __lt__(self, o):
        return self._bin(o, torch.lt)

    def __le__(self, o):
        return self._bin(o, torch.le)

    def __gt__(self, o):
        return self._bin(o, torch.gt)

    def __ge__(self, o):
        return self._bin(o, torch.ge)

    def __eq__(self, o):
        return self._bin(o, torch.eq)

    def __ne__(self, o):
        return self._bin(o, torch.ne)

    __hash__ = None


def unwrap_block(block) -> ArenaBlock:
    """The ``ArenaBlock`` inside a reference ``ArrayApiBlockBackend.BlockCls``.  The reference exposes no accessor for
    the wrapped object, but ``block.shape`` is answered by reading ``obj.shape`` (array_api.cpp Block::shape), so the
    wrapped object announces itself there; ``block[key]`` with a sentinel key (forwarded to ``obj.__getitem__``,
    array_api.cpp:181-189) is the slower fallback."""
    _LAST_SHAPE_READ[0] = None
    block.shape
    obj = _LAST_SHAPE_READ[0]
    if obj is not None:
        _LAST_SHAPE_READ[0] = None
        return obj
    key = _Unwrap()
    block[key]
    if key.obj is None:
        raise TypeError("block does not wrap an ArenaBlock (not produced by the cyten_b200 plugin's block backend)")
    return key.obj


class _Linalg:
    @staticmethod
    def eigh(a):
        w, v = torch.linalg.eigh(_t(a))
        return ArenaBlock(w), ArenaBlock(v)

    @staticmethod
    def eigvalsh(a):
        return ArenaBlock(torch.linalg.eigvalsh(_t(a)))

    @staticmethod
    def vector_norm(a, axis=None, ord=2):
        return ArenaBlock(torch.linalg.vector_norm(_t(a), ord=ord, dim=axis))

    @staticmethod
    def trace(a):
        return ArenaBlock(torch.diagonal(_t(a), dim1=-2, dim2=-1).sum(-1))      # batched over leading axes (Array API)

    @staticmethod
    def qr(a, mode="reduced"):
        q, r = torch.linalg.qr(_t(a), mode=mode)
        return ArenaBlock(q), ArenaBlock(r)

    @staticmethod
    def svd(a, full_matrices=False):
        u, s, vh = torch.linalg.svd(_t(a), full_matrices=full_matrices)
        return ArenaBlock(u), ArenaBlock(s), ArenaBlock(vh)


class TorchNamespace:
    """The subset of the Array API standard that ``ArrayApiBlockBackend`` calls, on torch tensors."""
    float32, float64 = torch.float32, torch.float64
    complex64, complex128 = torch.complex64, torch.complex128
    bool, int64 = torch.bool, torch.int64
    linalg = _Linalg()

    def __init__(self, default_device="cpu"):
        self.default_device = torch.device(default_device)

    def _dev(self, device):
        return self.default_device if device is None else torch.device(device)

    # -- creation -----------------------------------------------------------------------------------------------------
    def asarray(self, obj, dtype=None, device=None, copy=None):
        if type(obj) is ArenaBlock and dtype is None and device is None and not copy:
            return obj                                           # the hot case: as_block of an arena view
        if isinstance(obj, ArenaBlock):
            same_dtype = dtype is None or obj.dtype == dtype
            same_dev = device is None or torch.device(device) == torch.device(obj.device)
            if same_dtype and same_dev and not copy:
                return obj                                       # zero copy: the arena view itself
            t = obj.t
            if not same_dtype:
                t = self._cast(t, dtype)
            if not same_dev:
                t = t.to(torch.device(device))
            if copy and t is obj.t:
                t = t.clone()
            return ArenaBlock(t)
        if isinstance(obj, torch.Tensor):
            t = obj
        else:
            a = np.asarray(obj)
            t = _from_np(a) if a.dtype != object else torch.as_tensor(obj)
        if dtype is not None and t.dtype != dtype:
            t = self._cast(t, dtype)
        t = t.to(self._dev(device))
        if copy:
            t = t.clone()
        return ArenaBlock(t)

    @staticmethod
    def _cast(t, dtype):
        if t.is_complex() and not dtype.is_complex and dtype != torch.bool:
            t = t.real                                           # numpy's astype drops the imaginary part
        return t.to(dtype)

    def zeros(self, shape, dtype=None, device=None):
        return ArenaBlock(torch.zeros(tuple(shape), dtype=dtype or torch.float64, device=self._dev(device)))

    def ones(self, shape, dtype=None, device=None):
        return ArenaBlock(torch.ones(tuple(shape), dtype=dtype or torch.float64, device=self._dev(device)))

    def eye(self, n, m=None, dtype=None, device=None):
        return ArenaBlock(torch.eye(int(n), int(n if m is None else m), dtype=dtype or torch.float64, device=self._dev(device)))

    # -- element-wise -------------------------------------------------------------------------------------------------
    def abs(self, a):
        return ArenaBlock(torch.abs(_t(a)))

    def conj(self, a):
        return ArenaBlock(torch.conj_physical(_t(a)))

    def real(self, a):
        t = _t(a)
        return ArenaBlock(t.real.clone() if t.is_complex() else t)

    def imag(self, a):
        t = _t(a)
        return ArenaBlock(t.imag.clone() if t.is_complex() else torch.zeros_like(t))

    def exp(self, a):
        return ArenaBlock(torch.exp(_t(a)))

    def log(self, a):
        return ArenaBlock(torch.log(_t(a)))

    def sqrt(self, a):
        return ArenaBlock(torch.sqrt(_t(a)))

    def where(self, cond, a, b):
        c = _t(cond)
        a, b = _tt(a, c), _tt(b, c)
        if not isinstance(a, torch.Tensor):
            ref = b if isinstance(b, torch.Tensor) else None
            a = torch.as_tensor(a, dtype=ref.dtype if ref is not None else torch.float64, device=c.device)
        if not isinstance(b, torch.Tensor):
            b = torch.as_tensor(b, dtype=a.dtype, device=c.device)
        return ArenaBlock(torch.where(c, a, b))

    def astype(self, a, dtype, copy=True):
        return ArenaBlock(self._cast(_t(a), dtype))

    # -- reductions ---------------------------------------------------------------------------------------------------
    def sum(self, a, axis=None):
        t = _t(a)
        return ArenaBlock(torch.sum(t) if axis is None else torch.sum(t, dim=axis))

    def max(self, a, axis=None):
        t = _t(a)
        return ArenaBlock(torch.max(t) if axis is None else torch.amax(t, dim=axis))

    def min(self, a, axis=None):
        t = _t(a)
        return ArenaBlock(torch.min(t) if axis is None else torch.amin(t, dim=axis))

    def all(self, a):
        return ArenaBlock(torch.all(_t(a)))

    def any(self, a):
        return ArenaBlock(torch.any(_t(a)))

    def argmax(self, a):
        return ArenaBlock(torch.argmax(_t(a)))

    def argmin(self, a):
        return ArenaBlock(torch.argmin(_t(a)))

    def argsort(self, a, axis=-1):
        return ArenaBlock(torch.argsort(_t(a), dim=axis, stable=True))

    # -- shape manipulation -------------------------------------------------------------------------------------------
    def reshape(self, a, shape):
        return ArenaBlock(torch.reshape(_t(a), tuple(shape)))

    def permute_dims(self, a, axes):
        return ArenaBlock(_t(a).permute(*[int(x) for x in axes]))

    def expand_dims(self, a, axis=0):
        return ArenaBlock(torch.unsqueeze(_t(a), int(axis)))

    def squeeze(self, a, axis):
        t = _t(a)
        axes = (axis,) if isinstance(axis, int) else tuple(int(x) for x in axis)
        return ArenaBlock(torch.squeeze(t, dim=axes) if axes else t)

    def concat(self, arrays, axis=0):
        return ArenaBlock(torch.cat([_t(x) for x in arrays], dim=axis))

    def diagonal(self, a):
        return ArenaBlock(torch.diagonal(_t(a)).clone())

    # -- linear algebra -----------------------------------------------------------------------------------------------
    def matmul(self, a, b):
        return ArenaBlock(torch.matmul(_t(a), _t(b)))

    def tensordot(self, a, b, axes=2):
        ta, tb = _t(a), _t(b)
        if ta.dtype != tb.dtype:
            dt = torch.promote_types(ta.dtype, tb.dtype)
            ta, tb = ta.to(dt), tb.to(dt)
        if isinstance(axes, int):
            return ArenaBlock(torch.tensordot(ta, tb, dims=axes))
        return ArenaBlock(torch.tensordot(ta, tb, dims=([int(x) for x in axes[0]], [int(x) for x in axes[1]])))


NAMESPACE = TorchNamespace


def make_block_backend(cyten, device: str):
    """``ArrayApiBlockBackend`` over :class:`TorchNamespace`, with the handful of block operations the reference's
    generic Array API backend leaves unimplemented (src/block_backend/array_api.cpp:678, 782, 788, 838, 934, 956, 1002,
    1111) or implements as an endless ``mul`` <-> ``Block::operator*`` recursion (block_backend.cpp:153-156, 694-697:
    ``BlockBackend::mul(Scalar, Block)`` calls ``Block * Scalar``, which calls ``get_backend()->mul`` again; the numpy and
    torch backends override ``mul``, the Array API one does not).  The trampoline ``PyArrayApiBlockBackend``
    (pybind/block_backend/py_trampolines.hpp:496-862) makes these Python overrides visible to the C++ callers.
    Cold-path glue only."""
    base = cyten.block_backends.ArrayApiBlockBackend
    t2d = {torch.float32: cyten.Dtype.float32, torch.float64: cyten.Dtype.float64, torch.complex64: cyten.Dtype.complex64,
           torch.complex128: cyten.Dtype.complex128, torch.bool: cyten.Dtype.bool, torch.int64: cyten.Dtype.int64}
    d2t = {v: k for k, v in t2d.items()}

    class ArenaBlockBackend(base):
        def _w(self, t: torch.Tensor):
            """torch tensor -> Block of this backend, keeping its dtype."""
            return self.as_block(ArenaBlock(t), t2d.get(t.dtype))

        @staticmethod
        def _t(block) -> torch.Tensor:
            return unwrap_block(block).t

        @staticmethod
        def _num(a):
            if hasattr(a, "to_numpy"):
                a = a.to_numpy()
            if isinstance(a, np.ndarray):
                a = a.item()
            if isinstance(a, np.generic):
                a = a.item()
            return a

        def mul(self, a, b):
            return self._w(torch.mul(self._t(b), self._num(a)))

        def angle(self, a):
            return self._w(torch.angle(self._t(a)))

        def sqrt(self, a):
            return self._w(torch.sqrt(self._t(a)))

        def kron(self, a, b):
            return self._w(torch.kron(self._t(a), self._t(b)))

        def matrix_exp(self, matrix):
            return self._w(torch.linalg.matrix_exp(self._t(matrix)))

        def real_if_close(self, a, tol):
            t = self._t(a)
            if t.is_complex():
                eps = torch.finfo(t.real.dtype).eps
                if bool((t.imag.abs() <= tol * eps).all()):          # numpy: tol > 1 is in units of machine epsilon
                    t = t.real.clone()
            return self._w(t)

        def tile(self, a, repeats):
            t = self._t(a)
            return self._w(t.repeat(*([int(repeats)] if t.ndim <= 1 else [1] * (t.ndim - 1) + [int(repeats)])))

        def trace_partial(self, a, idcs1, idcs2, remaining_idcs):
            # the generic implementation flattens the remaining axes (array_api.cpp:1070-1086); the numpy / torch
            # backends keep them (numpy.cpp:1167-1195)
            t = self._t(a)
            rem = [int(i) % t.ndim for i in remaining_idcs]
            i1, i2 = [int(i) % t.ndim for i in idcs1], [int(i) % t.ndim for i in idcs2]
            t = t.permute(*rem, *i1, *i2)
            keep = list(t.shape[:len(rem)])
            d = 1
            for s_ in t.shape[len(rem):len(rem) + len(i1)]:
                d *= int(s_)
            t = t.reshape(*keep, d, d)
            return self._w(torch.diagonal(t, dim1=-2, dim2=-1).sum(-1))

        def block_from_diagonal(self, diag):
            return self._w(torch.diag(self._t(diag)))

        def block_from_mask(self, mask, dtype):
            m = self._t(mask).to(torch.bool)
            (M,) = m.shape
            cols = torch.nonzero(m).flatten()
            res = torch.zeros((len(cols), M), dtype=d2t[dtype], device=m.device)
            res[torch.arange(len(cols), device=m.device), cols] = 1
            return self._w(res)

    return ArenaBlockBackend(TorchNamespace(device), device)

"""Host side of the Krylov vector operations: ``inner``, ``norm``,
``linear_combination``, ``mul``, ``dagger``, ``scale_axis`` of the reference's ``AbelianBackend``
(src/backends/abelian.cpp:2159-2211, 2780-2791, 2253-2301, 2759-2778, 693-703, 3177-3232).

The reference merges the two block lists on the host and calls numpy once per block.  Here the merge is a
cached plan (a device segment table per pair of structures) and one launch covers every block; results that
a Krylov loop only feeds back into the next vector operation (alpha of the Lanczos recurrence, the norm of
the new vector) can stay on the device (``*_dev`` variants) so that a whole iteration is free of host
synchronisation and can be captured into a CUDA graph together with the matvec.
"""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import numpy as np
import torch

from . import _lib
from .data import AbelianBackendData, aligned_offsets, ptr, to_device_i64


def _row_keys(block_inds: np.ndarray):
    return [r.tobytes() for r in np.ascontiguousarray(block_inds, dtype=np.int64)]


class Blas1Mixin:
    """Mixed into ``B200AbelianBackend`` (uses its plan cache, ``_call``, ``_ctx``, ``_stream``, ``_empty``)."""

    # -- plumbing ---------------------------------------------------------------------------------------
    def _b1_lib(self):
        return self._library if self._library is not None else _lib.load()

    def _b1_workspace(self) -> torch.Tensor:
        """Reduction workspace (per-CTA partials + arrival counter, re-armed by the last CTA): ONE per (backend, stream).
        Two dot / lincomb-with-norm launches on different streams would otherwise share counters and partials; the same
        holds for a graph replay overlapping an eager call, so graphs are captured and replayed on the stream they run on."""
        import torch as _t
        key = _t.cuda.current_stream(self.device).cuda_stream if self.device.type == "cuda" else 0
        pool = self.__dict__.setdefault("_b1_work_by_stream", {})
        w = pool.get(key)
        if w is None:
            n = int(self._b1_lib().cyb_blas1_workspace_bytes())
            w = pool[key] = torch.zeros(n, dtype=torch.uint8, device=self.device)   # zeroed once, kernels re-arm it
        return w

    def _b1_workspace_single(self) -> torch.Tensor:
        w = getattr(self, "_b1_work", None)
        if w is None:
            n = int(self._b1_lib().cyb_blas1_workspace_bytes())
            w = self._b1_work = torch.zeros(n, dtype=torch.uint8, device=self.device)   # zeroed once, kernels re-arm it
        return w

    def _b1_table(self, rows: np.ndarray, dtype: str) -> dict:
        """rows int64[nseg][ncol], last column = block size -> device segment table + chunk prefix."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        chunk = int(self._b1_lib().cyb_blas1_chunk_elems(_lib.CYB_C128 if dtype == "complex128" else _lib.CYB_F64))
        nchunks = (rows[:, -1] + chunk - 1) // chunk if len(rows) else np.zeros(0, np.int64)
        chunk0 = np.zeros(len(rows) + 1, dtype=np.int64)
        chunk0[1:] = np.cumsum(nchunks)
        with self._ctx():
            return dict(seg=to_device_i64(rows if len(rows) else np.zeros((1, rows.shape[1])), self.device),
                        chunk0=to_device_i64(chunk0, self.device), nseg=len(rows), nchunk=int(chunk0[-1]))

    def _b1_pairs(self, a: AbelianBackendData, b: AbelianBackendData):
        """Positions (i, j) of the blocks present in both tensors (both lists are lexsorted the same way)."""
        if a.sig == b.sig or np.array_equal(a.block_inds, b.block_inds):
            idx = np.arange(a.num_blocks)
            return idx, idx
        where = {k: j for j, k in enumerate(_row_keys(b.block_inds))}
        ij = [(i, where[k]) for i, k in enumerate(_row_keys(a.block_inds)) if k in where]
        ij = np.asarray(ij, dtype=np.int64).reshape(-1, 2)
        return ij[:, 0], ij[:, 1]

    @staticmethod
    def _b1_common_dtype(*dtypes) -> str:
        return "complex128" if any(d == "complex128" for d in dtypes) else "float64"

    # -- inner / norm -----------------------------------------------------------------------------------
    def inner_dev(self, a: AbelianBackendData, b: AbelianBackendData, do_dagger: bool = True,
                  real_only: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
        """<a|b> (do_dagger) or sum_i a[i1..in] b[in..i1] as a device tensor float64[2] = (re, im); no host sync."""
        dt = self._b1_common_dtype(a.dtype, b.dtype)
        if a.dtype != dt:
            a = self._to_dtype(a, dt)
        if b.dtype != dt:
            b = self._to_dtype(b, dt)
        if not do_dagger:      # trace(a b): b's blocks transposed (all axes reversed), then matched like a's
            r = b.block_inds.shape[1]
            b = self.permute_legs(SimpleNamespace(data=b), list(reversed(range(r))), [])
        key = ("b1dot", a.sig, b.sig)
        plan = self._plan_get(key)
        if plan is None:
            i, j = self._b1_pairs(a, b)
            if len(i) and not np.array_equal(a.shapes[i], b.shapes[j]):
                raise ValueError("inner: the tensors have different legs")
            rows = np.stack([a.offsets[i], b.offsets[j], a.sizes[i]], axis=1) if len(i) else np.zeros((0, 3), np.int64)
            plan = self._plan_put(key, self._b1_table(rows, dt))
        with self._ctx():
            if out is None:
                out = torch.empty(2, dtype=torch.float64, device=self.device)
            if self.dry_run:
                return out
            self._call("cyb_block_dot", ptr(a.buf), ptr(b.buf), ptr(plan["seg"]), ptr(plan["chunk0"]), plan["nseg"],
                       plan["nchunk"], a.cyb_dtype, (1 if do_dagger else 0) | (2 if real_only else 0), ptr(out),
                       ptr(self._b1_workspace()), self._stream())
        return out

    def inner(self, a: AbelianBackendData, b: AbelianBackendData, do_dagger: bool = True):
        """abelian.cpp:2159-2211.  Python scalar (one 16-byte read-back)."""
        re, im = self.inner_dev(a, b, do_dagger).tolist()
        return complex(re, im) if "complex128" in (a.dtype, b.dtype) else re

    def norm_sq_dev(self, a: AbelianBackendData, out: torch.Tensor | None = None) -> torch.Tensor:
        return self.inner_dev(a, a, True, real_only=True, out=out)

    def norm(self, data: AbelianBackendData) -> float:
        """Frobenius norm over all blocks (abelian.cpp:2780-2791): one launch, the arena padding is skipped."""
        return float(np.sqrt(self.norm_sq_dev(data)[0].item()))

    # -- linear_combination / mul -----------------------------------------------------------------------
    def _b1_lincomb(self, ops, coefs, res: AbelianBackendData, plan, coef_dev=None, sumsq=None):
        coef = np.zeros(6, dtype=np.float64)
        for i, c in enumerate(coefs):
            coef[2 * i], coef[2 * i + 1] = np.real(c), np.imag(c)
        dev = None
        if coef_dev is not None:
            dev = (C.c_void_p * 3)(*[(None if t is None else t.data_ptr()) for t in (list(coef_dev) + [None] * 3)[:3]])
        bufs = [(None if o is None else o.buf) for o in (list(ops) + [None] * 3)[:3]]
        with self._ctx():
            if self.dry_run:
                return res
            self._call("cyb_block_lincomb", ptr(bufs[0]), ptr(bufs[1]), ptr(bufs[2]), ptr(res.buf), ptr(plan["seg"]),
                       ptr(plan["chunk0"]), plan["nseg"], plan["nchunk"], res.cyb_dtype,
                       coef.ctypes.data_as(C.c_void_p), dev, ptr(sumsq),
                       ptr(self._b1_workspace()) if sumsq is not None else None, self._stream())
        return res

    def linear_combination(self, a, v: AbelianBackendData, b, w: AbelianBackendData) -> AbelianBackendData:
        """abelian.cpp:2253-2301: a*v + b*w on the union of the blocks."""
        dt = self._b1_common_dtype(v.dtype, w.dtype, *("complex128" for c in (a, b) if np.iscomplexobj(c)))
        v = v if v.dtype == dt else self._to_dtype(v, dt)
        w = w if w.dtype == dt else self._to_dtype(w, dt)
        key = ("b1lin", v.sig, w.sig)
        plan = self._plan_get(key)
        if plan is None:
            if v.sig == w.sig or np.array_equal(v.block_inds, w.block_inds):
                bi, shapes = v.block_inds, v.shapes
                ov, ow = v.offsets, w.offsets
            else:
                vk, wk = _row_keys(v.block_inds), _row_keys(w.block_inds)
                wpos = {k: j for j, k in enumerate(wk)}
                vset = set(vk)
                rows = [(v.block_inds[i], v.shapes[i], v.offsets[i], w.offsets[wpos[k]] if k in wpos else -1)
                        for i, k in enumerate(vk)]
                rows += [(w.block_inds[j], w.shapes[j], -1, w.offsets[j]) for j, k in enumerate(wk) if k not in vset]
                bi = np.asarray([r[0] for r in rows], dtype=np.int64).reshape(len(rows), v.block_inds.shape[1])
                order = self.lexsort(bi) if len(bi) > 1 else np.arange(len(bi))
                bi = bi[order]
                shapes = np.asarray([rows[t][1] for t in order], dtype=np.int64).reshape(len(rows), -1)
                ov = np.asarray([rows[t][2] for t in order], dtype=np.int64)
                ow = np.asarray([rows[t][3] for t in order], dtype=np.int64)
            sizes = np.prod(shapes, axis=1).astype(np.int64) if shapes.shape[1] else np.ones(len(shapes), np.int64)
            offs, _ = aligned_offsets(sizes)
            rows = np.stack([ov, ow, np.full(len(ov), -1), offs, sizes], axis=1) if len(ov) else np.zeros((0, 5), np.int64)
            plan = self._plan_put(key, dict(self._b1_table(rows, dt), bi=bi, shapes=shapes, sig=None))
        res = self._empty(dt, self.device, plan["bi"], plan["shapes"])
        if plan["sig"] is None:
            plan["sig"] = res.sig
        res._sig = plan["sig"]
        return self._b1_lincomb([v, w, None], [a, b, 0.0], res, plan)

    def lincomb3_dev(self, ops, coefs, coef_dev=None, sumsq: torch.Tensor | None = None,
                     out: AbelianBackendData | None = None) -> AbelianBackendData:
        """out = c0*x + c1*y + c2*z for up to three tensors of IDENTICAL structure; ``coef_dev[i]`` (device
        float64[2]) multiplies ``coefs[i]``; ``sumsq`` (device float64[1]) receives ||out||^2 in the same launch.
        The fused three-term recurrence + norm of a Lanczos step (krylov_based.cpp:856-873)."""
        first = next(o for o in ops if o is not None)
        for o in ops:
            if o is not None and (o.sig != first.sig):
                raise ValueError("lincomb3_dev needs tensors of identical structure (use linear_combination)")
        key = ("b1lin3", first.sig)
        plan = self._plan_get(key)
        if plan is None:
            off, sizes = first.offsets, first.sizes
            rows = np.stack([off, off, off, off, sizes], axis=1) if len(off) else np.zeros((0, 5), np.int64)
            plan = self._plan_put(key, self._b1_table(rows, first.dtype))
        if out is None:
            out = AbelianBackendData(first.dtype, self.device, self._new_buf(first.dtype, first.buf.numel()),
                                     first.block_inds, first.offsets, first.shapes, sig=first._sig)
        out._relayout = None                     # written in place: cached re-layouts of the old contents are stale
        return self._b1_lincomb(ops, coefs, out, plan, coef_dev=coef_dev, sumsq=sumsq)

    def mul(self, a, b: AbelianBackendData) -> AbelianBackendData:
        """abelian.cpp:2759-2778: a == 0 gives the tensor without blocks."""
        dt = "complex128" if np.iscomplexobj(a) else b.dtype
        if a == 0:
            r = b.block_inds.shape[1]
            return self._empty(dt, self.device, np.zeros((0, r), np.int64), np.zeros((0, r), np.int64))
        b = b if b.dtype == dt else self._to_dtype(b, dt)
        return self.lincomb3_dev([b, None, None], [a, 0.0, 0.0])

    # -- dagger -------------------------------------------------------------------------------------------
    def dagger(self, a: AbelianBackendData) -> AbelianBackendData:
        """abelian.cpp:693-703: every block conjugate-transposed (all axes reversed), block_inds columns reversed
        and re-sorted.  Real tensors stay a deferred permutation; complex ones are one conjugating block copy."""
        r = a.block_inds.shape[1]
        res = self.permute_legs(SimpleNamespace(data=a), list(reversed(range(r))), [])
        if a.dtype != "complex128" or self.dry_run:
            return res
        if not res.is_pending:                 # r <= 1: same arena, conjugate into a new one
            return self.lincomb_conj(res)
        src, axes_perm, order = res.pending_source
        plan = self._plans.get(("permute", src.sig, tuple(axes_perm))) if self.plan_cache else None
        if plan is None:                       # plan cache off: rebuild the job list
            new_shapes = src.shapes[:, list(axes_perm)][order]
            from .backend import _c_strides
            jobs = self._make_jobs(src.offsets[order], res.offsets, new_shapes,
                                   _c_strides(src.shapes)[:, list(axes_perm)][order], _c_strides(new_shapes))
            copy = self._prepare_copy(jobs)
        else:
            copy = plan["copy"]
        buf = self._new_buf(a.dtype, int(res.offsets[-1] + res.sizes[-1]) if res.num_blocks else 0)
        self._run_copy(src.buf, buf, copy, src.cyb_dtype, conj=True)
        out = AbelianBackendData(a.dtype, self.device, buf, res.block_inds, res.offsets, res.shapes, sig=res._sig)
        return out

    def lincomb_conj(self, a: AbelianBackendData) -> AbelianBackendData:
        """Elementwise complex conjugate into a new arena (rank-1 dagger, DiagonalTensor dagger)."""
        from .backend import _c_strides
        st = _c_strides(a.shapes)
        jobs = self._make_jobs(a.offsets, a.offsets, a.shapes, st, st)
        buf = self._new_buf(a.dtype, a.buf.numel())
        self._copy_blocks(a.buf, buf, jobs, a.cyb_dtype, conj=(a.dtype == "complex128"))
        return AbelianBackendData(a.dtype, self.device, buf, a.block_inds, a.offsets, a.shapes, sig=a._sig)

    # -- scale_axis -----------------------------------------------------------------------------------------
    def scale_axis(self, a: AbelianBackendData, d: AbelianBackendData, leg: int) -> AbelianBackendData:
        """abelian.cpp:3177-3232: block (.., i_leg, ..) times the diagonal block of sector i_leg along axis ``leg``;
        blocks whose sector has no diagonal block drop out."""
        dt = self._b1_common_dtype(a.dtype, d.dtype)
        a = a if a.dtype == dt else self._to_dtype(a, dt)
        key = ("b1scale", a.sig, d.sig, int(leg))
        plan = self._plan_get(key)
        if plan is None:
            dpos = {int(s): j for j, s in enumerate(d.block_inds[:, 0])} if d.num_blocks else {}
            keep = np.asarray([i for i in range(a.num_blocks) if int(a.block_inds[i, leg]) in dpos], dtype=np.int64)
            bi, shapes = a.block_inds[keep], a.shapes[keep]
            sizes = np.prod(shapes, axis=1).astype(np.int64) if len(keep) else np.zeros(0, np.int64)
            offs, _ = aligned_offsets(sizes)
            desc = np.zeros((len(keep), 6), dtype=np.int64)
            for t, i in enumerate(keep):
                sh = a.shapes[i]
                desc[t] = (a.offsets[i], offs[t], int(np.prod(sh[:leg])), sh[leg], int(np.prod(sh[leg + 1:])),
                           d.offsets[dpos[int(a.block_inds[i, leg])]])
                if d.sizes[dpos[int(a.block_inds[i, leg])]] != sh[leg]:
                    raise ValueError("scale_axis: the diagonal tensor's leg does not match the tensor's leg")
            work0 = np.zeros(len(keep) + 1, dtype=np.int64)
            chunk = int(self._b1_lib().cyb_blas1_chunk_elems(_lib.CYB_C128 if dt == "complex128" else _lib.CYB_F64))
            work0[1:] = np.cumsum((sizes + chunk - 1) // chunk)
            with self._ctx():
                plan = self._plan_put(key, dict(bi=bi, shapes=shapes, sig=None, nblk=len(keep), nwork=int(work0[-1]),
                                                desc=to_device_i64(desc if len(keep) else np.zeros((1, 6)), self.device),
                                                work0=to_device_i64(work0, self.device)))
        res = self._empty(dt, self.device, plan["bi"], plan["shapes"])
        if plan["sig"] is None:
            plan["sig"] = res.sig
        res._sig = plan["sig"]
        if self.dry_run or plan["nblk"] == 0:
            return res
        with self._ctx():
            self._call("cyb_scale_axis", ptr(a.buf), ptr(res.buf), ptr(plan["desc"]), ptr(plan["work0"]), plan["nblk"],
                       plan["nwork"], ptr(d.buf), a.cyb_dtype, d.cyb_dtype, self._stream())
        return res

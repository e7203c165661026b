"""``B200AbelianBackend``: the hot-path operations of the reference's ``AbelianBackend``, driven from
Python, executed by the CUDA kernels behind ``libcyten_b200.so``.

Method names, argument meaning and result conventions follow ``cyten::AbelianBackend``
(src/backends/abelian.cpp; interface include/cyten/backends/tensor_backend.h:85-421):
``permute_legs`` (:988-1006), ``combine_legs`` (:1022-1219), ``split_legs`` (:3235-3436),
``compose`` (:1222-1515), ``svd`` (:3461-3568), ``qr`` (:3084-3151), ``eigh`` (:1759-1788),
``truncate_singular_values`` (:3623-3638).  A tensor here is anything with ``.data``
(:class:`AbelianBackendData`), ``.codomain`` / ``.domain`` (:class:`TensorProduct`) -- see
``cyten_b200.tensors``.

There is no CPU path: every data-touching step is a kernel launch through the C-ABI.
"""
from __future__ import annotations

import contextlib
import ctypes as C

import numpy as np
import torch

from . import _lib
from .blas1 import Blas1Mixin
from .data import AbelianBackendData, aligned_offsets, ptr, stream_ptr, to_device_i64, ALIGN
from .spaces import AbelianLegPipe, ElementarySpace, TensorProduct

MAX_RANK = _lib.MAX_RANK

# eigh(sort=...): BlockBackend::argsort (block_backend.cpp:759-781) applied to REAL eigenvalues that the kernels
# deliver in ascending order.  None / '<' / 'SR' / 'SA' leave them as they are; the imaginary-part orders sort all-zero
# keys; '>' reverses; 'm<' / 'm>' sort by magnitude.
_EIGH_SORTS = {None: None, "<": None, "SR": None, "SA": None, "SI": "zero", "LI": "-zero",
               ">": "rev", "LR": "rev", "LA": "rev", "m<": "abs", "SM": "abs", "m>": "-abs", "LM": "-abs"}


def conventional_leg_order(codomain: TensorProduct, domain: TensorProduct):
    """``[*codomain, *reversed(domain)]``: the axis order of every block."""
    return [*codomain.factors, *reversed(domain.factors)]


def _c_strides(shapes: np.ndarray) -> np.ndarray:
    """Row-major element strides for a batch of shapes (n, r)."""
    n, r = shapes.shape
    st = np.ones((n, r), dtype=np.int64)
    for a in range(r - 2, -1, -1):
        st[:, a] = st[:, a + 1] * shapes[:, a + 1]
    return st


class B200AbelianBackend(Blas1Mixin):
    """Abelian block-sparse backend on one B200."""

    def __init__(self, device="cuda:0", _library=None):
        """``_library`` is a test hook: an object with the C-ABI's functions (``call(name, *args)``),
        used by the CPU test-suite to drive this host logic with the numpy oracle.  The product path
        never passes it: the CUDA library is mandatory and the device must be a GPU."""
        self.device = torch.device(device)
        self._library = _library
        if _library is None:
            if self.device.type != "cuda":
                raise _lib.CytenB200Error("B200AbelianBackend needs a CUDA device (no CPU fallback)")
            _lib.load()
            self._lib_call = _lib.call
            self._plan_ws = lambda na, nb: _lib.load().cyb_compose_plan_workspace(na, nb)
            self._copy_ws = lambda jobs, n: _lib.load().cyb_copy_blocks_workspace(jobs.ctypes.data_as(C.c_void_p), n)
            self._gemm_ws = lambda n: _lib.load().cyb_grouped_gemm_workspace(n)
            self._gemm_skinny_rows = int(_lib.load().cyb_grouped_gemm_skinny_rows())
        else:
            self._lib_call = _library.call
            self._plan_ws = _library.cyb_compose_plan_workspace
            self._copy_ws = lambda jobs, n: _library.cyb_copy_blocks_workspace(jobs.ctypes.data_as(C.c_void_p), n)
            self._gemm_ws = _library.cyb_grouped_gemm_workspace
            self._gemm_skinny_rows = int(_library.cyb_grouped_gemm_skinny_rows())
        # Plans: the integer bookkeeping of an operation (sorted block_inds, copy job lists, GEMM work
        # lists and tile schedules) depends only on the structure of its operands (block_inds + legs),
        # not on the block contents.  It is computed ON DEVICE on the first call and replayed for every
        # later call with the same structure (the Lanczos / sweep loops of DMRG call the same
        # contraction thousands of times); a replay is pure kernel launches with no host<->device sync.
        self.plan_cache = True
        self._plans = {}
        self.plan_hits = self.plan_misses = 0
        self.launches = 0   # number of cyten_b200 kernel-launching C-ABI calls (bench bookkeeping)
        self.last_info = None
        self.dry_run = False        # plan-only mode: tables are computed, arenas are not touched
        self.plan_mkn = None        # bench hook: list collecting sum(M*K*N) of every compose
        self.gemm_events = None     # bench hook: list of (start, end) CUDA events around the GEMM launch
        self.gemm_event_flops = []
        self.shard = None           # multi-GPU hook, see cyten_b200.sharding
        self.check_convergence = True   # raise when a Jacobi factorisation hits its sweep limit

    def _empty(self, dtype, device, block_inds, shapes, zero=False):
        if self.dry_run:
            block_inds = np.asarray(block_inds, dtype=np.int64)
            shapes = np.asarray(shapes, dtype=np.int64)
            shapes = shapes.reshape(len(block_inds), -1) if len(block_inds) else shapes.reshape(0, block_inds.shape[1])
            sizes = np.prod(shapes, axis=1) if shapes.shape[1] else np.ones(len(shapes), np.int64)
            offs, _ = aligned_offsets(sizes)
            return AbelianBackendData(dtype, device, torch.empty(ALIGN, dtype=torch.float64), block_inds, offs, shapes)
        return AbelianBackendData.empty(dtype, device, block_inds, shapes, zero=zero)

    def _ctx(self):
        return torch.cuda.device(self.device) if self.device.type == "cuda" else contextlib.nullcontext()

    def _stream(self):
        return stream_ptr(self.device) if self.device.type == "cuda" else C.c_void_p(0)

    # ------------------------------------------------------------------------------------------
    # small device-bookkeeping wrappers (a1)
    # ------------------------------------------------------------------------------------------
    # entry points that only read / write integer tables (the plan-only mode may run them)
    _TABLE_ONLY = frozenset(("cyb_lexsort_rows", "cyb_row_differences", "cyb_pack_rows", "cyb_take_rows", "cyb_common_sorted",
                             "cyb_fuse_charges", "cyb_compose_plan", "cyb_copy_blocks_prepare", "cyb_grouped_gemm_prepare"))

    def _call(self, name, *args):
        if self.dry_run and name not in self._TABLE_ONLY:
            # arenas are 16-element dummies in plan-only mode: a kernel on them would read / write out of bounds
            raise _lib.CytenB200Error(f"{name}: not available in plan-only mode (dry_run covers permute_legs / combine_legs / "
                                      "split_legs / compose and their bookkeeping, not factorisations or reductions)")
        self.launches += 1
        self._lib_call(name, *args)

    def lexsort(self, rows: np.ndarray) -> np.ndarray:
        """Device ``np.lexsort(rows.T)`` (BlockInds::lexsort_indices, block_inds.cpp:290-299)."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        n, r = rows.shape
        if n <= 1 or r == 0:
            return np.arange(n, dtype=np.int64)
        with self._ctx():
            d_rows = to_device_i64(rows, self.device)
            d_perm = torch.empty(n, dtype=torch.int64, device=self.device)
            self._call("cyb_lexsort_rows", ptr(d_rows), n, r, ptr(d_perm), self._stream())
            return d_perm.cpu().numpy()

    def row_differences(self, rows: np.ndarray) -> np.ndarray:
        """Device ``find_row_differences(include_len=True)`` (block_inds.cpp:308-336)."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        n, r = rows.shape
        if n == 0:
            return np.zeros(1, dtype=np.int64)
        with self._ctx():
            d_rows = to_device_i64(rows, self.device)
            d_out = torch.empty(n + 2, dtype=torch.int64, device=self.device)
            self._call("cyb_row_differences", ptr(d_rows), n, r, ptr(d_out),
                       C.c_void_p(d_out.data_ptr() + 8 * (n + 1)), self._stream())
            out = d_out.cpu().numpy()
        return out[:int(out[n + 1])].copy()

    def _copy_blocks(self, src_buf, dst_buf, jobs: np.ndarray, cyb_dtype: int, conj: bool = False):
        """jobs: int64 (njobs, 27) rows = cyb_copy_job."""
        njobs = len(jobs)
        if njobs == 0 or self.dry_run:
            return
        with self._ctx():
            jobs = np.ascontiguousarray(jobs, dtype=np.int64)
            d_jobs = to_device_i64(jobs, self.device)
            self._call("cyb_copy_blocks", ptr(src_buf), ptr(dst_buf), ptr(d_jobs), jobs.ctypes.data_as(C.c_void_p),
                       njobs, cyb_dtype, int(conj), self._stream())

    def _prepare_copy(self, jobs: np.ndarray):
        """Uploads a job list and builds its device schedule once; returns (njobs, workspace)."""
        njobs = len(jobs)
        if njobs == 0 or self.dry_run:
            return (0, None)
        with self._ctx():
            jobs = np.ascontiguousarray(jobs, dtype=np.int64)
            d_jobs = to_device_i64(jobs, self.device)
            nbytes = int(self._copy_ws(jobs, njobs))
            d_ws = torch.empty(nbytes // 8 + 1, dtype=torch.int64, device=self.device)
            self._call("cyb_copy_blocks_prepare", ptr(d_jobs), njobs, ptr(d_ws), nbytes, self._stream())
        return (njobs, d_ws)

    def _run_copy(self, src_buf, dst_buf, prepared, cyb_dtype: int, conj: bool = False):
        njobs, d_ws = prepared
        if njobs == 0 or self.dry_run:
            return
        with self._ctx():
            self._call("cyb_copy_blocks_run", ptr(src_buf), ptr(dst_buf), ptr(d_ws), njobs, cyb_dtype, int(conj),
                       self._stream())

    def _plan_get(self, key):
        if not self.plan_cache:
            return None
        plan = self._plans.get(key)
        if plan is None:
            self.plan_misses += 1
        else:
            self.plan_hits += 1
        return plan

    def _plan_put(self, key, plan):
        # plan-only (dry-run) passes build tables without device schedules / workspaces: never cache those
        if self.plan_cache and not self.dry_run:
            self._plans[key] = plan
        return plan

    def _new_buf(self, dtype: str, total: int) -> torch.Tensor:
        tdt = torch.complex128 if dtype == "complex128" else torch.float64
        if self.dry_run:
            t = torch.empty(ALIGN, dtype=tdt)
            t._cyb_dummy = True        # must never be cached as a tensor's arena (AbelianBackendData.buf)
            return t
        return torch.empty(max(int(total), ALIGN), dtype=tdt, device=self.device)

    @staticmethod
    def _make_jobs(src_off, dst_off, shapes, src_strides, dst_strides) -> np.ndarray:
        n, r = shapes.shape
        if r > MAX_RANK:
            raise _lib.CytenB200Error(f"blocks of rank {r} > {MAX_RANK} are not supported")
        jobs = np.zeros((n, _lib.COPY_JOB_WORDS), dtype=np.int64)
        jobs[:, 0] = src_off
        jobs[:, 1] = dst_off
        jobs[:, 2] = r
        jobs[:, 3:3 + r] = shapes
        jobs[:, 3 + MAX_RANK:3 + MAX_RANK + r] = src_strides
        jobs[:, 3 + 2 * MAX_RANK:3 + 2 * MAX_RANK + r] = dst_strides
        return jobs

    # ------------------------------------------------------------------------------------------
    # (a3) permute_legs
    # ------------------------------------------------------------------------------------------
    def permute_legs(self, a, codomain_idcs, domain_idcs, new_codomain=None, new_domain=None,
                     mixes_codomain_domain=None, levels=None, bend_right=None) -> AbelianBackendData:
        """abelian.cpp:988-1006: permute the axes of every block, the columns of block_inds, re-sort.

        The integer part (column permutation + device lexsort + job list) is a cached plan; the data
        movement is deferred until a consumer asks for the arena (the reference's numpy blocks are lazy
        views at the same place), so that ``compose`` can fold it into its own operand re-layout.  A
        permutation that leaves the block axis order unchanged (legs only bent between codomain and
        domain) moves no data at all."""
        data: AbelianBackendData = a.data
        axes_perm = [int(i) for i in codomain_idcs] + [int(i) for i in reversed(list(domain_idcs))]
        r = data.block_inds.shape[1]
        if sorted(axes_perm) != list(range(r)):
            raise ValueError("codomain_idcs + domain_idcs must be a permutation of the legs")
        if axes_perm == list(range(r)):
            if data.is_pending:
                data.buf   # materialise once, the alias shares the arena
            res = AbelianBackendData(data.dtype, self.device, data.buf, data.block_inds, data.offsets,
                                     data.shapes, sig=data._sig)
            res._dev_tables = data._dev_tables
            return res
        key = ("permute", data.sig, tuple(axes_perm))
        plan = self._plan_get(key)
        if plan is None:
            new_bi = data.block_inds[:, axes_perm]
            order = self.lexsort(new_bi)
            new_bi = new_bi[order]
            old_shapes = data.shapes
            new_shapes = old_shapes[:, axes_perm][order]
            sizes = np.prod(new_shapes, axis=1).astype(np.int64) if r else np.ones(len(new_shapes), np.int64)
            offs, total = aligned_offsets(sizes)
            jobs = self._make_jobs(data.offsets[order], offs, new_shapes,
                                   _c_strides(old_shapes)[:, axes_perm][order], _c_strides(new_shapes))
            plan = self._plan_put(key, dict(bi=new_bi, shapes=new_shapes, offsets=offs, total=total, order=order,
                                            copy=self._prepare_copy(jobs), sig=None))
        backend = self

        def materialise():
            buf = backend._new_buf(data.dtype, plan["total"])
            backend._run_copy(data.buf, buf, plan["copy"], data.cyb_dtype)
            return buf
        res = AbelianBackendData(data.dtype, self.device, None, plan["bi"], plan["offsets"], plan["shapes"],
                                 sig=plan["sig"], pending=(materialise, (data, tuple(axes_perm), plan["order"])))
        if plan["sig"] is None:
            plan["sig"] = res.sig
        return res

    # ------------------------------------------------------------------------------------------
    # (a4) combine_legs
    # ------------------------------------------------------------------------------------------
    def combine_legs(self, tensor, leg_idcs_combine, pipes, new_codomain: TensorProduct,
                     new_domain: TensorProduct) -> AbelianBackendData:
        """abelian.cpp:1022-1219.  ``leg_idcs_combine``: list of contiguous ascending index groups,
        ``pipes[g]``: the AbelianLegPipe of group g (as it appears in new_codomain / new_domain)."""
        data: AbelianBackendData = tensor.data
        num_legs = data.block_inds.shape[1]
        num_codomain = tensor.codomain.num_factors
        old_legs = conventional_leg_order(tensor.codomain, tensor.domain)
        groups = [[int(i) for i in g] for g in leg_idcs_combine]
        for p in pipes:
            if not isinstance(p, AbelianLegPipe):
                raise ValueError("abelian backend requires AbelianLegPipe")
        num_result_legs = num_legs - sum(len(g) - 1 for g in groups)
        # the integer part below depends on the structure only: cached plan (tables + prepared copy schedule)
        key = ("combine", data.sig, num_codomain, tuple(tuple(g) for g in groups), tuple(p.sig for p in pipes),
               new_codomain.sig, new_domain.sig)
        plan = self._plan_get(key)
        if plan is not None:
            res = self._empty(data.dtype, self.device, plan["bi"], plan["shapes"], zero=True)
            res._sig = plan["sig"]
            self._run_copy(data.buf, res.buf, plan["copy"], data.cyb_dtype)
            return res
        t_bi = data.block_inds
        n_old = data.num_blocks
        res_bi = np.zeros((n_old, num_result_legs), dtype=np.int64)
        map_rows = []       # per group: row of pipe.block_ind_map for every old block
        new_axis_of_group = []
        old_to_new_axis = np.zeros(num_legs, dtype=np.int64)
        cstyles = []
        i = j = 0
        for g, pipe in zip(groups, pipes):
            while j < g[0]:
                res_bi[:, i] = t_bi[:, j]
                old_to_new_axis[j] = i
                i += 1
                j += 1
            in_domain = g[0] >= num_codomain
            cstyles.append(pipe.combine_cstyle != in_domain)
            bi_group = t_bi[:, g[0]:g[-1] + 1]
            if in_domain:
                bi_group = bi_group[:, ::-1]
            multi = bi_group @ pipe.sector_strides
            rows = pipe._inv_fusion_outcomes_sort[multi]
            map_rows.append(rows)
            res_bi[:, i] = pipe.block_ind_map[rows, -1]
            new_axis_of_group.append(i)
            for k in g:
                old_to_new_axis[k] = i
            i += 1
            j += len(g)
        while j < num_legs:
            res_bi[:, i] = t_bi[:, j]
            old_to_new_axis[j] = i
            i += 1
            j += 1
        # sort (duplicates stay adjacent), then dedupe -> new blocks
        sort = self.lexsort(res_bi)
        res_bi = res_bi[sort]
        map_rows = [m[sort] for m in map_rows]
        diffs = self.row_differences(res_bi)
        new_bi = res_bi[diffs[:-1]]
        new_legs = conventional_leg_order(new_codomain, new_domain)
        new_shapes = np.stack([np.asarray(l.multiplicities)[new_bi[:, ax]] for ax, l in enumerate(new_legs)],
                              axis=1) if num_result_legs else np.zeros((len(new_bi), 0), np.int64)
        res = self._empty(data.dtype, self.device, new_bi, new_shapes, zero=True)
        if n_old == 0:
            return res
        # every old block -> a box inside its new block
        new_block_of_old = np.repeat(np.arange(len(new_bi)), np.diff(diffs))
        old_shapes = data.shapes[sort]
        new_strides = _c_strides(new_shapes)[new_block_of_old]          # (n_old, num_result_legs)
        dst_strides = np.zeros((n_old, num_legs), dtype=np.int64)
        dst_off = res.offsets[new_block_of_old].copy()
        in_group = np.zeros(num_legs, dtype=bool)
        for g, pipe, rows, cst, ax in zip(groups, pipes, map_rows, cstyles, new_axis_of_group):
            in_group[g] = True
            gshape = old_shapes[:, g[0]:g[-1] + 1]
            # position inside the combined axis: C-order over the group's axes (F-order if !cstyle)
            inner = np.ones_like(gshape)
            if cst:
                for k in range(len(g) - 2, -1, -1):
                    inner[:, k] = inner[:, k + 1] * gshape[:, k + 1]
            else:
                for k in range(1, len(g)):
                    inner[:, k] = inner[:, k - 1] * gshape[:, k - 1]
            dst_strides[:, g[0]:g[-1] + 1] = inner * new_strides[:, ax:ax + 1]
            dst_off += pipe.block_ind_map[rows, 0] * new_strides[:, ax]
        for k in range(num_legs):
            if not in_group[k]:
                dst_strides[:, k] = new_strides[:, old_to_new_axis[k]]
        jobs = self._make_jobs(data.offsets[sort], dst_off, old_shapes, _c_strides(old_shapes), dst_strides)
        prepared = self._prepare_copy(jobs)
        self._plan_put(key, dict(bi=new_bi, shapes=new_shapes, copy=prepared, sig=res.sig))
        self._run_copy(data.buf, res.buf, prepared, data.cyb_dtype)
        return res

    # ------------------------------------------------------------------------------------------
    # (a5) split_legs
    # ------------------------------------------------------------------------------------------
    def split_legs(self, a, leg_idcs, new_codomain: TensorProduct, new_domain: TensorProduct
                   ) -> AbelianBackendData:
        """abelian.cpp:3235-3436: every old block is cut into the boxes listed by the pipes'
        block_ind_map; each box becomes one new block (reshaped, F-style groups axis-reversed)."""
        data: AbelianBackendData = a.data
        a_num_legs = data.block_inds.shape[1]
        num_codomain = a.codomain.num_factors
        old_legs = conventional_leg_order(a.codomain, a.domain)
        leg_idcs = [int(i) for i in leg_idcs]
        pipes = [old_legs[i] for i in leg_idcs]
        for p in pipes:
            if not isinstance(p, AbelianLegPipe):
                raise ValueError("split_legs: leg is not an AbelianLegPipe")
        new_legs = conventional_leg_order(new_codomain, new_domain)
        res_num_legs = len(new_legs)
        n_old = data.num_blocks
        if n_old == 0:
            return self._empty(data.dtype, self.device, np.zeros((0, res_num_legs), np.int64),
                                            np.zeros((0, res_num_legs), np.int64))
        key = ("split", data.sig, num_codomain, tuple(leg_idcs), tuple(p.sig for p in pipes),
               new_codomain.sig, new_domain.sig)
        plan = self._plan_get(key)
        if plan is not None:
            res = self._empty(data.dtype, self.device, plan["bi"], plan["shapes"])
            res._sig = plan["sig"]
            self._run_copy(data.buf, res.buf, plan["copy"], data.cyb_dtype)
            return res
        old_bi = data.block_inds
        n_split = len(leg_idcs)
        beg = np.stack([p.block_ind_map_slices[old_bi[:, i]] for p, i in zip(pipes, leg_idcs)], axis=1)
        cnt = np.stack([p.block_ind_map_slices[old_bi[:, i] + 1] for p, i in zip(pipes, leg_idcs)], axis=1) - beg
        per_old = np.prod(cnt, axis=1)
        old_rows = np.repeat(np.arange(n_old), per_old)
        n_new = len(old_rows)
        # map_rows[:, j]: row of pipes[j].block_ind_map; C-order enumeration inside each old block
        start = np.concatenate([[0], np.cumsum(per_old)])[:-1]
        local = np.arange(n_new) - np.repeat(start, per_old)
        map_rows = np.zeros((n_new, n_split), dtype=np.int64)
        cnt_rep = cnt[old_rows]
        rem = local.copy()
        for jj in range(n_split - 1, -1, -1):
            map_rows[:, jj] = rem % cnt_rep[:, jj] + beg[old_rows, jj]
            rem //= cnt_rep[:, jj]
        new_bi = np.zeros((n_new, res_num_legs), dtype=np.int64)
        src_beg = np.zeros((n_new, a_num_legs), dtype=np.int64)
        # per NEW axis: which old axis it lives in and its stride multiplier inside that axis
        new_axis_old = np.zeros(res_num_legs, dtype=np.int64)
        groups_new = []
        shift = 0
        jp = 0
        for i_leg in range(a_num_legs):
            if i_leg in leg_idcs:
                pipe = pipes[jp]
                in_domain = i_leg >= num_codomain
                k, k2 = i_leg + shift, i_leg + shift + pipe.num_legs
                bmap = pipe.block_ind_map[map_rows[:, jp]]
                cols = bmap[:, 2:2 + pipe.num_legs]
                new_bi[:, k:k2] = cols[:, ::-1] if in_domain else cols
                src_beg[:, i_leg] = bmap[:, 0]
                new_axis_old[k:k2] = i_leg
                groups_new.append((k, k2, pipe.combine_cstyle != in_domain))
                shift += pipe.num_legs - 1
                jp += 1
            else:
                new_bi[:, i_leg + shift] = old_bi[old_rows, i_leg]
                new_axis_old[i_leg + shift] = i_leg
        new_shapes = np.stack([np.asarray(l.multiplicities)[new_bi[:, ax]] for ax, l in enumerate(new_legs)], axis=1)
        order = self.lexsort(new_bi)
        new_bi, new_shapes, old_rows, src_beg = new_bi[order], new_shapes[order], old_rows[order], src_beg[order]
        res = self._empty(data.dtype, self.device, new_bi, new_shapes)
        old_strides = _c_strides(data.shapes)[old_rows]                 # (n_new, a_num_legs)
        src_strides = old_strides[:, new_axis_old].copy()               # (n_new, res_num_legs)
        for (k, k2, cst) in groups_new:
            gshape = new_shapes[:, k:k2]
            inner = np.ones_like(gshape)
            if cst:
                for q in range(k2 - k - 2, -1, -1):
                    inner[:, q] = inner[:, q + 1] * gshape[:, q + 1]
            else:
                for q in range(1, k2 - k):
                    inner[:, q] = inner[:, q - 1] * gshape[:, q - 1]
            src_strides[:, k:k2] *= inner
        src_off = data.offsets[old_rows] + np.sum(src_beg * old_strides, axis=1)
        jobs = self._make_jobs(src_off, res.offsets, new_shapes, src_strides, _c_strides(new_shapes))
        prepared = self._prepare_copy(jobs)
        self._plan_put(key, dict(bi=new_bi, shapes=new_shapes, copy=prepared, sig=res.sig))
        self._run_copy(data.buf, res.buf, prepared, data.cyb_dtype)
        return res

    # ------------------------------------------------------------------------------------------
    # (a6) compose
    # ------------------------------------------------------------------------------------------
    def _fuse_charges(self, d_bi, nrows, ncols, cols, legs, symmetry):
        """Device fused charge of the legs ``legs`` (block_inds columns ``cols``) per row."""
        nsym = symmetry.sector_ind_len
        out = torch.zeros((max(nrows, 1), nsym), dtype=torch.int64, device=self.device)
        if len(legs) == 0 or nrows == 0:
            return out
        sectors = np.concatenate([l.sector_decomposition for l in legs], axis=0)
        sector_ptr = np.concatenate([[0], np.cumsum([l.num_sectors for l in legs])])[:-1]
        d_cols = to_device_i64(np.asarray(cols), self.device)
        d_sec = to_device_i64(sectors, self.device)
        d_sptr = to_device_i64(sector_ptr, self.device)
        d_mod = to_device_i64(symmetry.moduli_array(), self.device)
        self._call("cyb_fuse_charges", ptr(d_bi), nrows, ncols, ptr(d_cols), len(legs), ptr(d_sec),
                   ptr(d_sptr), ptr(d_mod), nsym, ptr(out), self._stream())
        return out

    def compose_plan(self, a, b):
        """Device planner of abelian_compose_worker (abelian.cpp:1238-1469): returns the host copies
        of the result tables plus the device work list for :meth:`_run_gemm`."""
        A: AbelianBackendData = a.data
        B: AbelianBackendData = b.data
        a_keep = a.codomain.num_factors
        ncontr = a.domain.num_factors
        if ncontr != b.codomain.num_factors:
            raise ValueError("compose: a.domain and b.codomain have different numbers of legs")
        b_keep = b.domain.num_factors
        na, nb = A.num_blocks, B.num_blocks
        ra, rb = a_keep + ncontr, ncontr + b_keep
        sym = a.codomain.symmetry
        dev = self.device
        a_shapes, b_shapes = A.shapes, B.shapes
        a_mk = np.stack([np.prod(a_shapes[:, :a_keep], axis=1), np.prod(a_shapes[:, a_keep:], axis=1)], axis=1)
        b_kn = np.stack([np.prod(b_shapes[:, :ncontr], axis=1), np.prod(b_shapes[:, ncontr:], axis=1)], axis=1)
        contr_nsec = np.asarray([l.num_sectors for l in b.codomain.factors], dtype=np.int64)
        with self._ctx():
            d_abi, d_bbi = A.dev("block_inds"), B.dev("block_inds")
            # coupled charges: codomain legs of a (columns 0..a_keep-1); domain legs of b: factor f
            # sits in column rb-1-f (abelian.cpp:1384-1418)
            qa = self._fuse_charges(d_abi, na, ra, list(range(a_keep)), a.codomain.factors, sym)
            qb = self._fuse_charges(d_bbi, nb, rb, [rb - 1 - f for f in range(b_keep)], b.domain.factors, sym)
            d_nsec, d_amk, d_bkn = (to_device_i64(contr_nsec, dev), to_device_i64(a_mk, dev),
                                    to_device_i64(b_kn, dev))
            ws_bytes = self._plan_ws(na, nb)
            d_ws = torch.empty(ws_bytes // 8 + 1, dtype=torch.int64, device=dev)
            rr = a_keep + b_keep
            cap_out = min(na * nb, 8 * (na + nb) + 64)
            cap_pairs = min(na * nb, 32 * (na + nb) + 64)
            for attempt in range(2):
                d_tab = torch.empty(8 + cap_out * (rr + 5) + (cap_out + 1) + 3 * cap_pairs,
                                    dtype=torch.int64, device=dev)
                o = 0

                def take(n):
                    nonlocal o
                    t = d_tab[o:o + n]
                    o += n
                    return t
                d_header, d_out_bi, d_rowcol = take(8), take(cap_out * rr), take(cap_out * 2)
                d_mn, d_off, d_pp, d_pairs = take(cap_out * 2), take(cap_out), take(cap_out + 1), take(3 * cap_pairs)
                self._call("cyb_compose_plan", ptr(d_abi), na, ra, a_keep, ptr(d_bbi), nb, rb, ncontr,
                           ptr(d_nsec), ptr(qa), ptr(qb), sym.sector_ind_len, ptr(d_amk), ptr(d_bkn),
                           ptr(A.dev("offsets")), ptr(B.dev("offsets")), ALIGN, cap_out, cap_pairs,
                           ptr(d_header), ptr(d_out_bi), ptr(d_rowcol), ptr(d_mn), ptr(d_off), ptr(d_pp),
                           ptr(d_pairs), ptr(d_ws), ws_bytes, self._stream())
                header = d_header.cpu().numpy()
                if header[3]:
                    break
                cap_out, cap_pairs = int(header[0]) + 1, int(header[1]) + 1   # exact counts, retry once
            else:
                raise _lib.CytenB200Error("compose_plan: table capacity exceeded "
                                          f"({header[0]} blocks, {header[1]} pairs)")
            n_out = int(header[0])
            out_bi = d_out_bi[:n_out * rr].cpu().numpy().reshape(n_out, rr)
            out_mn = d_mn[:n_out * 2].cpu().numpy().reshape(n_out, 2)
            out_off = d_off[:n_out].cpu().numpy()
            rowcol = d_rowcol[:n_out * 2].cpu().numpy().reshape(n_out, 2)
        plan = dict(n_out=n_out, n_pairs=int(header[1]), total_elems=int(header[2]), mkn=int(header[4]),
                    out_bi=out_bi, out_mn=out_mn, out_off=out_off, rowcol=rowcol,
                    d_mn=d_mn, d_off=d_off, d_pp=d_pp, d_pairs=d_pairs, _keep=(d_tab,))
        return plan

    def compose(self, a, b, subset=None) -> AbelianBackendData:
        """abelian.cpp:1222-1234 + 1238-1469: block-sparse matrix product over the matched
        contracted sectors = (cached) device plan + ONE grouped GEMM call (DMMA kernel for the blocks
        with M > 8, streaming kernel for the skinny ones)."""
        A: AbelianBackendData = a.data
        B: AbelianBackendData = b.data
        a_keep, ncontr, b_keep = a.codomain.num_factors, a.domain.num_factors, b.domain.num_factors
        if ncontr == 0:
            return self._compose_no_contraction(a, b)
        res_dtype = "complex128" if "complex128" in (A.dtype, B.dtype) else "float64"
        rr = a_keep + b_keep
        if A.num_blocks == 0 or B.num_blocks == 0:
            return self._empty(res_dtype, self.device, np.zeros((0, rr), np.int64),
                                            np.zeros((0, rr), np.int64))
        if A.dtype != res_dtype:
            a = _Retyped(a, self._to_dtype(A, res_dtype))
            A = a.data
        if B.dtype != res_dtype:
            b = _Retyped(b, self._to_dtype(B, res_dtype))
            B = b.data
        sub_key = None if subset is None else tuple(int(x) for x in subset)
        key = ("compose", A.sig, B.sig, a.codomain.sig, a.domain.sig, b.domain.sig, sub_key)
        plan = self._plan_get(key)
        if plan is None:
            plan = self.compose_plan(a, b)
            n_out = plan["n_out"]
            legs = conventional_leg_order(a.codomain, b.domain)
            out_bi = plan["out_bi"]
            plan["shapes"] = (np.stack([np.asarray(l.multiplicities)[out_bi[:, ax]] for ax, l in enumerate(legs)], axis=1)
                              if rr else np.zeros((n_out, 0), np.int64))
            plan["sig"] = None
            plan["nrun"] = n_out if subset is None else len(subset)
            plan["gemm_ws"] = None
            # which of the two kernels have work (a launch without work costs ~13 us): streaming kernel for blocks of
            # at most SK rows, DMMA kernel for the others
            mn = plan["out_mn"] if subset is None else plan["out_mn"][np.asarray(subset, dtype=np.int64)]
            live = (mn[:, 0] > 0) & (mn[:, 1] > 0)
            plan["gemm_kernels"] = (1 if np.any(live & (mn[:, 0] <= self._gemm_skinny_rows)) else 0) | \
                                   (2 if np.any(live & (mn[:, 0] > self._gemm_skinny_rows)) else 0)
            if plan["nrun"] and not self.dry_run:
                with self._ctx():
                    d_sub = None if subset is None else to_device_i64(np.asarray(subset), self.device)
                    nbytes = int(self._gemm_ws(plan["nrun"]))
                    ws = torch.empty(nbytes // 8 + 1, dtype=torch.int64, device=self.device)
                    self._call("cyb_grouped_gemm_prepare", _lib.CYB_C128 if res_dtype == "complex128" else _lib.CYB_F64,
                               ptr(plan["d_mn"]), ptr(plan["d_pp"]), ptr(plan["d_pairs"]), ptr(d_sub), plan["nrun"],
                               ptr(ws), nbytes, self._stream())
                    plan["gemm_ws"] = ws
            self._plan_put(key, plan)
        if self.plan_mkn is not None:
            self.plan_mkn.append(plan["mkn"])
        # A's trailing axes are the contracted legs in REVERSED order; bring B's leading axes into the
        # same order (abelian.cpp:1362-1381) -- one batched copy, fused with a pending permute_legs of B
        b_buf = self._compose_operand_b(B, ncontr, b_keep) if ncontr > 1 else B.buf
        buf = self._new_buf(res_dtype, plan["total_elems"])
        res = AbelianBackendData(res_dtype, self.device, buf, plan["out_bi"], plan["out_off"], plan["shapes"],
                                 sig=plan["sig"])
        if plan["sig"] is None:
            plan["sig"] = res.sig
        self._run_gemm(A.buf, b_buf, res, plan)
        res._dev_tables["offsets"] = plan["d_off"][:plan["n_out"]]
        return res

    def _compose_no_contraction(self, a, b) -> AbelianBackendData:
        """abelian.cpp:1518-1565: a has no domain, b no codomain -- every pair of blocks gives an outer product.
        That is the ordinary compose over ONE extra leg holding only the trivial sector with multiplicity 1 (K = 1):
        the tables get one more column, the arenas are shared, the planner and the grouped GEMM run unchanged."""
        sym = a.codomain.symmetry
        triv = ElementarySpace(sym, np.zeros((1, sym.sector_ind_len), dtype=np.int64), np.ones(1, dtype=np.int64), False)

        def with_leg(data, front):
            z, o = np.zeros((data.num_blocks, 1), np.int64), np.ones((data.num_blocks, 1), np.int64)
            bi = np.concatenate([z, data.block_inds] if front else [data.block_inds, z], axis=1)
            shapes = np.concatenate([o, data.shapes] if front else [data.shapes, o], axis=1)
            return AbelianBackendData(data.dtype, self.device, data.buf, bi, data.offsets, shapes)
        a2 = _Retyped(a, with_leg(a.data, False))
        a2.domain = TensorProduct([triv], symmetry=sym)
        b2 = _Retyped(b, with_leg(b.data, True))
        b2.codomain = TensorProduct([triv], symmetry=sym)
        return self.compose(a2, b2)

    def _compose_operand_b(self, B: AbelianBackendData, ncontr: int, b_keep: int) -> torch.Tensor:
        """Arena with every block of B re-laid-out to axes ``[reversed(contracted), kept]`` (same block
        offsets).  If B is a not-yet-materialised permute_legs of some parent tensor, the permutation
        and the reversal are done in ONE pass straight from the parent's arena."""
        perm = list(range(ncontr - 1, -1, -1)) + list(range(ncontr, ncontr + b_keep))
        src = B.pending_source
        if src is not None:
            parent, axes_perm, order = src
            key = ("operand_b", parent.sig, axes_perm, ncontr)
            plan = self._plan_get(key)
            if plan is None:
                shp = B.shapes[:, perm]
                src_strides = _c_strides(parent.shapes)[order][:, list(axes_perm)][:, perm]
                jobs = self._make_jobs(parent.offsets[order], B.offsets, shp, src_strides, _c_strides(shp))
                plan = self._plan_put(key, dict(copy=self._prepare_copy(jobs)))
            src_buf = parent.buf
        else:
            # an operand that is reused (the environments / MPO of a Lanczos loop) keeps its re-laid-out copy
            if B._relayout is not None and ncontr in B._relayout:
                return B._relayout[ncontr]
            key = ("operand_b", B.sig, None, ncontr)
            plan = self._plan_get(key)
            if plan is None:
                shp = B.shapes[:, perm]
                jobs = self._make_jobs(B.offsets, B.offsets, shp, _c_strides(B.shapes)[:, perm], _c_strides(shp))
                plan = self._plan_put(key, dict(copy=self._prepare_copy(jobs)))
            src_buf = B.buf
        total = int(B.offsets[-1] + B.sizes[-1]) if B.num_blocks else 0
        out = self._new_buf(B.dtype, total)
        self._run_copy(src_buf, out, plan["copy"], B.cyb_dtype)
        if src is None and not self.dry_run:
            if B._relayout is None:
                B._relayout = {}
            B._relayout[ncontr] = out
        return out

    def _run_gemm(self, a_buf, b_buf, res, plan):
        if plan["nrun"] == 0 or self.dry_run:
            return
        timed = self.gemm_events is not None and self.device.type == "cuda"
        if timed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        with self._ctx():
            self._call("cyb_grouped_gemm_run", ptr(a_buf), ptr(b_buf), ptr(res.buf), res.cyb_dtype,
                       ptr(plan["d_mn"]), ptr(plan["d_off"]), ptr(plan["d_pp"]), ptr(plan["d_pairs"]),
                       plan["nrun"], ptr(plan["gemm_ws"]), plan["gemm_kernels"], self._stream())
        if timed:
            e1.record()
            self.gemm_events.append((e0, e1))
            self.gemm_event_flops.append((8.0 if res.dtype == "complex128" else 2.0) * plan["mkn"])

    def _to_dtype(self, data: AbelianBackendData, dtype: str) -> AbelianBackendData:
        tdt = torch.complex128 if dtype == "complex128" else torch.float64
        return AbelianBackendData(dtype, self.device, data.buf.to(tdt), data.block_inds, data.offsets, data.shapes)

    # ------------------------------------------------------------------------------------------
    # (a7, a9, a10) per-sector factorisations
    # ------------------------------------------------------------------------------------------
    def _matched_sectors(self, a, new_leg):
        """The merge-walk over common coupled sectors shared by svd / qr (abelian.cpp:3488-3541):
        yields (j, n, k, block index or -1): j/k index cod0/dom0, n indexes new_leg."""
        cod0, dom0 = a.codomain.factors[0], a.domain.factors[0]
        sc, sd = a.codomain.sector_decomposition, a.domain.sector_decomposition
        lut_d = {tuple(s): k for k, s in enumerate(sd.tolist())}
        where = {(int(r[0]), int(r[1])): i for i, r in enumerate(a.data.block_inds)}
        out = []
        n_enum = 0
        for j, sec in enumerate(sc.tolist()):
            k = lut_d.get(tuple(sec))
            if k is None:
                continue
            n = n_enum
            n_enum += 1
            jj = j if cod0.sector_order == "sorted" else cod0.sector_decomposition_where(sec)
            kk = k if dom0.sector_order == "sorted" else dom0.sector_decomposition_where(sec)
            nn = n if new_leg.sector_order == "sorted" else new_leg.sector_decomposition_where(sec)
            out.append((jj, nn, kk, where.get((jj, kk), -1)))
        return out

    def _sorted_data(self, dtype, block_inds, shapes):
        """Allocates an arena for the given blocks in lexsorted order; returns (data, position of
        every input row in the arena)."""
        block_inds, shapes = np.asarray(block_inds, dtype=np.int64), np.asarray(shapes, dtype=np.int64)
        if len(shapes) == 0:                       # no matched sector at all: tensors without blocks
            ncol = block_inds.shape[1] if block_inds.ndim == 2 else 2
            block_inds = np.zeros((0, ncol), np.int64)
            shapes = np.zeros((0, shapes.shape[1] if shapes.ndim == 2 else ncol), np.int64)
        else:
            block_inds = block_inds.reshape(len(shapes), -1)
            shapes = shapes.reshape(len(block_inds), -1)
        order = self.lexsort(block_inds) if len(block_inds) else np.zeros(0, np.int64)
        inv = np.empty(len(order), dtype=np.int64)
        inv[order] = np.arange(len(order))
        return self._empty(dtype, self.device, block_inds[order], shapes[order]), inv

    def _fz_out(self, dtype, block_inds, shapes, eye_rows=()):
        """Plan entry for one output of a factorisation: the lexsorted structure (``_sorted_data``), the position of every
        input row in it and the device tables of the blocks that are identity slices (absent sectors)."""
        data, pos = self._sorted_data(dtype, block_inds, shapes)
        eye = pos[np.asarray(eye_rows, dtype=np.int64)] if len(eye_rows) else np.zeros(0, np.int64)
        ent = dict(bi=data.block_inds, shapes=data.shapes, sig=data.sig, pos=pos, offsets=data.offsets, neye=len(eye), eye=None)
        if len(eye) and not self.dry_run:
            with self._ctx():
                ent["eye"] = (to_device_i64(data.shapes[eye], self.device), to_device_i64(data.offsets[eye], self.device))
        return ent

    def _fz_new(self, dtype, ent) -> AbelianBackendData:
        """A fresh arena with the cached structure; identity slices written by one launch from cached device tables."""
        data = self._empty(dtype, self.device, ent["bi"], ent["shapes"])
        data._sig = ent["sig"]
        if ent["neye"] and not self.dry_run:
            with self._ctx():
                self._call("cyb_fill_eye", ptr(data.buf), ptr(ent["eye"][0]), ptr(ent["eye"][1]), ent["neye"], data.cyb_dtype,
                           self._stream())
        return data

    def _fill_eye(self, data: AbelianBackendData, which):
        which = np.asarray(which, dtype=np.int64)
        if len(which) == 0 or self.dry_run:
            return
        with self._ctx():
            d_dims = to_device_i64(data.shapes[which], self.device)
            d_off = to_device_i64(data.offsets[which], self.device)
            self._call("cyb_fill_eye", ptr(data.buf), ptr(d_dims), ptr(d_off), len(which), data.cyb_dtype,
                       self._stream())

    def svd(self, a, new_co_domain: TensorProduct, algorithm=None, only_rows=None):
        """abelian.cpp:3461-3568: thin SVD of every present coupled sector in ONE batched launch;
        absent sectors get identity slices in U / Vh and no S block.  Returns (U, S, Vh) data.
        ``only_rows`` (multi-GPU sharding): positions in the matched-sector walk this rank owns; the
        other sectors are left to their owners (no block here)."""
        A: AbelianBackendData = a.data
        new_leg = new_co_domain.factors[0]
        # The sector walk, the three output structures (three device lexsorts with a read-back each), the identity-slice
        # tables and the descriptor table depend on the STRUCTURE of ``a`` only: a cached plan like every other operation
        # (the host bookkeeping was 1-4 ms of every svd / qr call).
        key = ("svd", A.sig, a.codomain.sig, a.domain.sig, new_leg.sig, None if only_rows is None else tuple(int(t) for t in only_rows))
        plan = self._plan_get(key)
        if plan is None:
            cod0, dom0 = a.codomain.factors[0], a.domain.factors[0]
            rows = self._matched_sectors(a, new_leg)
            if only_rows is not None:
                rows = [rows[t] for t in only_rows]
            u_bi = [(j, n) for j, n, k, i in rows]
            vh_bi = [(n, k) for j, n, k, i in rows]
            u_shapes = [(cod0.multiplicities[j], new_leg.multiplicities[n]) for j, n, k, i in rows]
            vh_shapes = [(new_leg.multiplicities[n], dom0.multiplicities[k]) for j, n, k, i in rows]
            present = [t for t, r in enumerate(rows) if r[3] >= 0]
            absent = [t for t, r in enumerate(rows) if r[3] < 0]
            s_bi = [(rows[t][1], rows[t][1]) for t in present]
            s_shapes = [(new_leg.multiplicities[rows[t][1]],) for t in present]
            pu = self._fz_out(A.dtype, np.asarray(u_bi).reshape(-1, 2), np.asarray(u_shapes).reshape(-1, 2), absent)
            pv = self._fz_out(A.dtype, np.asarray(vh_bi).reshape(-1, 2), np.asarray(vh_shapes).reshape(-1, 2), absent)
            ps = self._fz_out("float64", np.asarray(s_bi).reshape(-1, 2), np.asarray(s_shapes).reshape(-1, 1))
            ps["bi"] = np.asarray(ps["bi"]).reshape(-1, 2)
            desc = np.zeros((len(present), 6), dtype=np.int64)
            for q, t in enumerate(present):
                i = rows[t][3]
                m, n = A.shapes[i]
                if min(m, n) != ps["shapes"][ps["pos"][q], 0]:
                    raise _lib.CytenB200Error("svd: new leg multiplicity != min(m, n)")
                desc[q] = (m, n, A.offsets[i], pu["offsets"][pu["pos"][t]], ps["offsets"][ps["pos"][q]], pv["offsets"][pv["pos"][t]])
            plan = self._plan_put(key, dict(U=pu, Vh=pv, S=ps, desc=desc))
        U, Vh, S = self._fz_new(A.dtype, plan["U"]), self._fz_new(A.dtype, plan["Vh"]), self._fz_new("float64", plan["S"])
        S.block_inds = plan["S"]["bi"]
        desc = plan["desc"]
        info = None
        if len(desc):
            with self._ctx():
                info = torch.zeros(len(desc), dtype=torch.int64, device=self.device)
                self._call("cyb_batched_svd", ptr(A.buf), ptr(U.buf), ptr(S.buf), ptr(Vh.buf),
                           desc.ctypes.data_as(C.c_void_p), len(desc), A.cyb_dtype, ptr(info), self._stream())
        self.last_info = info
        self._check_converged(info, "svd")
        return U, S, Vh

    def qr(self, a, new_co_domain: TensorProduct):
        """abelian.cpp:3084-3151: thin QR per present sector (batched); identity slices for Q of
        absent sectors, no R block there."""
        A: AbelianBackendData = a.data
        new_leg = new_co_domain.factors[0]
        key = ("qr", A.sig, a.codomain.sig, a.domain.sig, new_leg.sig)
        plan = self._plan_get(key)
        if plan is None:                    # structure-only bookkeeping, cached (see svd)
            cod0, dom0 = a.codomain.factors[0], a.domain.factors[0]
            rows = self._matched_sectors(a, new_leg)
            present = [t for t, r in enumerate(rows) if r[3] >= 0]
            absent = [t for t, r in enumerate(rows) if r[3] < 0]
            q_bi = np.asarray([(j, n) for j, n, k, i in rows]).reshape(-1, 2)
            q_shapes = np.asarray([(cod0.multiplicities[j], new_leg.multiplicities[n]) for j, n, k, i in rows]).reshape(-1, 2)
            r_bi = np.asarray([(rows[t][1], rows[t][2]) for t in present]).reshape(-1, 2)
            r_shapes = np.asarray([(new_leg.multiplicities[rows[t][1]], dom0.multiplicities[rows[t][2]]) for t in present]).reshape(-1, 2)
            pq = self._fz_out(A.dtype, q_bi, q_shapes, absent)
            pr = self._fz_out(A.dtype, r_bi, r_shapes)
            desc = np.zeros((len(present), 6), dtype=np.int64)
            for q, t in enumerate(present):
                i = rows[t][3]
                m, n = A.shapes[i]
                desc[q] = (m, n, A.offsets[i], pq["offsets"][pq["pos"][t]], pr["offsets"][pr["pos"][q]], 0)
            plan = self._plan_put(key, dict(Q=pq, R=pr, desc=desc))
        Q, R = self._fz_new(A.dtype, plan["Q"]), self._fz_new(A.dtype, plan["R"])
        desc = plan["desc"]
        if len(desc):
            with self._ctx():
                self._call("cyb_batched_qr", ptr(A.buf), ptr(Q.buf), ptr(R.buf),
                           desc.ctypes.data_as(C.c_void_p), len(desc), A.cyb_dtype, self._stream())
        return Q, R

    def eigh(self, a, new_leg_dual: bool, sort=None):
        """abelian.cpp:1759-1788: per-sector Hermitian eigendecomposition; V starts as eye_data
        (abelian.cpp:1790-1826) and is overwritten where ``a`` has blocks.  Returns (W, V, new_leg)."""
        if sort not in _EIGH_SORTS:
            raise ValueError(f"Unknown sort option: '{sort}'")      # BlockBackend::argsort, block_backend.cpp:759-781
        A: AbelianBackendData = a.data
        dom0 = a.domain.factors[0]
        new_leg = a.domain.as_ElementarySpace(new_leg_dual)
        key = ("eigh", A.sig, a.domain.sig)
        plan = self._plan_get(key)
        if plan is None:                    # structure-only bookkeeping, cached (see svd)
            nsec = dom0.num_sectors
            v_bi = np.stack([np.arange(nsec), np.arange(nsec)], axis=1).astype(np.int64)
            v_shapes = np.stack([dom0.multiplicities, dom0.multiplicities], axis=1)
            has = np.zeros(nsec, dtype=bool)
            has[A.block_inds[:, 0]] = True
            pv = self._fz_out(A.dtype, v_bi, v_shapes, np.nonzero(~has)[0])      # v_bi is sorted already: pos = identity
            Wl = self._empty("float64", self.device, A.block_inds, A.shapes[:, :1])
            desc = np.zeros((A.num_blocks, 6), dtype=np.int64)
            for i in range(A.num_blocks):
                m, n = A.shapes[i]
                desc[i] = (m, n, A.offsets[i], 0, Wl.offsets[i], pv["offsets"][pv["pos"][A.block_inds[i, 0]]])
            plan = self._plan_put(key, dict(V=pv, w_shapes=A.shapes[:, :1].copy(), w_sig=Wl.sig, desc=desc))
        V = self._fz_new(A.dtype, plan["V"])
        W = self._empty("float64", self.device, A.block_inds, plan["w_shapes"])
        W._sig = plan["w_sig"]
        desc = plan["desc"]
        info = None
        if A.num_blocks:
            with self._ctx():
                info = torch.zeros(A.num_blocks, dtype=torch.int64, device=self.device)
                self._call("cyb_batched_eigh", ptr(A.buf), ptr(W.buf), ptr(V.buf),
                           desc.ctypes.data_as(C.c_void_p), A.num_blocks, A.cyb_dtype, ptr(info), self._stream())
        self.last_info = info
        self._check_converged(info, "eigh")
        if _EIGH_SORTS[sort] is not None and A.num_blocks:
            W, V = self._eigh_reorder(A, W, V, _EIGH_SORTS[sort])
        return W, V, new_leg

    def _eigh_reorder(self, A, W, V, mode):
        """``perm = argsort(w, sort); w = take(w, perm); v = take(v, perm, axis=1)`` per sector
        (NumpyBlockBackend::eigh, numpy.cpp:658-680).  The kernels deliver ascending eigenvalues: descending order
        is a reversal (no read-back); the two by-magnitude orders read the eigenvalues back once (dim(leg) doubles)
        to build the index lists.  The reordering itself is one gather launch for W and one for V."""
        nb = A.num_blocks
        dims = A.shapes[:, 0]
        if mode == "rev":
            idx_lists = [np.arange(int(d) - 1, -1, -1, dtype=np.int64) for d in dims]
        elif mode in ("zero", "-zero"):      # keys = imag(w) = 0: whatever numpy's argsort makes of equal keys
            idx_lists = [np.argsort(np.zeros(int(d)) * (1.0 if mode == "zero" else -1.0)).astype(np.int64) for d in dims]
        else:
            w_host = W.buf.cpu().numpy()
            idx_lists = []
            for i in range(nb):
                w = np.abs(w_host[W.offsets[i]:W.offsets[i] + dims[i]])
                idx_lists.append(np.argsort(w if mode == "abs" else -w).astype(np.int64))
        idx_ptr = np.concatenate([[0], np.cumsum(dims)]).astype(np.int64)
        W2 = AbelianBackendData(W.dtype, self.device, self._new_buf(W.dtype, W.buf.numel()), W.block_inds, W.offsets,
                                W.shapes, sig=W._sig)
        V2 = AbelianBackendData(V.dtype, self.device, V.buf.clone(), V.block_inds, V.offsets, V.shapes, sig=V._sig)
        dw = np.zeros((nb, 6), dtype=np.int64)
        dv = np.zeros((nb, 6), dtype=np.int64)
        for i in range(nb):
            vo = V.offsets[A.block_inds[i, 0]]
            dw[i] = (dims[i], 1, W.offsets[i], W.offsets[i], idx_ptr[i], dims[i])
            dv[i] = (dims[i], dims[i], vo, vo, idx_ptr[i], dims[i])
        with self._ctx():
            d_idx = to_device_i64(np.concatenate(idx_lists), self.device)
            self._call("cyb_gather_rows_cols", ptr(W.buf), ptr(W2.buf), ptr(to_device_i64(dw, self.device)), ptr(d_idx),
                       nb, 0, W.cyb_dtype, self._stream())
            self._call("cyb_gather_rows_cols", ptr(V.buf), ptr(V2.buf), ptr(to_device_i64(dv, self.device)), ptr(d_idx),
                       nb, 1, V.cyb_dtype, self._stream())
        return W2, V2

    def _check_converged(self, info, what: str):
        """A negative ``info`` entry = the Jacobi sweeps of that sector hit the sweep limit (factorize.cu); the
        reference's LAPACK path raises ``LinAlgError`` there, so this raises instead of returning unconverged
        factors.  One read-back of nsectors int64 (the factorisation has synchronised already)."""
        if info is None or not self.check_convergence or self.dry_run:
            return
        bad = (info < 0).nonzero().flatten().cpu().numpy() if hasattr(info, "nonzero") else np.nonzero(np.asarray(info) < 0)[0]
        if len(bad):
            raise _lib.CytenB200Error(f"{what}: Jacobi sweeps did not converge for sector(s) {bad.tolist()} "
                                      f"(CYB_ERR_NOCONV)")

    # ------------------------------------------------------------------------------------------
    # (a8) truncation
    # ------------------------------------------------------------------------------------------
    def diagonal_to_dense(self, S: AbelianBackendData, leg: ElementarySpace, scale: float = 1.0) -> torch.Tensor:
        """diagonal_tensor_to_block (abelian.cpp:1678-1692): zeros(leg.dim) with every block written
        at its sector's slice; optionally scaled."""
        dense = torch.zeros(max(leg.dim, 1), dtype=torch.float64, device=self.device)
        if S.num_blocks:
            sl = leg.slices[S.block_inds[:, 0], 0]
            shapes = S.shapes[:, :1]
            jobs = self._make_jobs(S.offsets, sl, shapes, np.ones_like(shapes), np.ones_like(shapes))
            self._copy_blocks(S.buf, dense, jobs, _lib.CYB_F64)
            if scale != 1.0:
                self._call("cyb_axpby", ptr(dense), ptr(dense), leg.dim, _lib.CYB_F64, float(scale), 0.0, self._stream())
        return dense

    def truncate_singular_values(self, S: AbelianBackendData, leg: ElementarySpace, chi_max=None,
                                 chi_min=1, degeneracy_tol=0.0, trunc_cut=0.0, svd_min=None,
                                 minimize_error=True, scale: float = 1.0):
        """abelian.cpp:3623-3638 + tensor_backend.cpp:138-242 + mask_from_block (abelian.cpp:2584-2639).

        Returns (keep mask over the leg's basis [host bool], small_leg, err, new_norm)."""
        dense = self.diagonal_to_dense(S, leg, scale)
        return self._select(dense, leg, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error)

    def truncate_dense(self, dense: torch.Tensor, leg: ElementarySpace, chi_max=None, chi_min=1,
                       degeneracy_tol=0.0, trunc_cut=0.0, svd_min=None, minimize_error=True):
        """Truncation of the NORMALISED singular values given as the dense vector over ``leg``
        (decompositions.cpp:672-709: ``S / norm(S)`` goes into truncate_singular_values).  Used where the
        vector was assembled by the multi-GPU all_gather.  Returns (keep, small_leg, err, new_norm, S_norm)."""
        n = leg.dim
        with self._ctx():
            out = torch.zeros(1, dtype=torch.float64, device=self.device)
            self._call("cyb_sumsq", ptr(dense), n, _lib.CYB_F64, ptr(out), self._stream())
            S_norm = float(out.sqrt().item())
            if S_norm > 0:
                self._call("cyb_axpby", ptr(dense), ptr(dense), n, _lib.CYB_F64, 1.0 / S_norm, 0.0, self._stream())
        return (*self._select(dense, leg, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error), S_norm)

    def cutoff_inverse(self, d: AbelianBackendData, cutoff: float) -> AbelianBackendData:
        """Pseudo-inverse of a real diagonal tensor: ``1/x`` where ``|x| >= cutoff``, else 0 (cyten.pinv of a
        DiagonalTensor, ops_algebra.cpp:1514-1515 -> BlockBackend::cutoff_inverse, numpy.cpp:645-656).  One launch over
        the whole arena (the padding between blocks is never read back)."""
        if d.dtype != "float64":
            raise _lib.CytenB200Error("cutoff_inverse: real diagonal tensors only")
        res = AbelianBackendData(d.dtype, self.device, self._new_buf(d.dtype, d.buf.numel()), d.block_inds, d.offsets,
                                 d.shapes, sig=d._sig)
        if d.num_blocks and not self.dry_run:
            n = int(d.offsets[-1] + d.sizes[-1])
            with self._ctx():
                self._call("cyb_cutoff_inverse", ptr(d.buf), ptr(res.buf), n, float(cutoff), self._stream())
        return res

    def scale_inplace(self, data: AbelianBackendData, factor: float):
        self._call("cyb_axpby", ptr(data.buf), ptr(data.buf), data.buf.numel(), data.cyb_dtype, float(factor), 0.0,
                   self._stream())

    def _select(self, dense, leg, chi_max, chi_min, degeneracy_tol, trunc_cut, svd_min, minimize_error):
        n = leg.dim
        with self._ctx():
            keep = torch.zeros(max(n, 1), dtype=torch.uint8, device=self.device)
            result = torch.zeros(4, dtype=torch.float64, device=self.device)
            self._call("cyb_truncate_select", ptr(dense), n, -1 if chi_max is None else int(chi_max),
                       int(chi_min), float(degeneracy_tol), float(trunc_cut),
                       -1.0 if svd_min is None else float(svd_min), int(bool(minimize_error)),
                       ptr(keep), ptr(result), self._stream())
            keep_h = keep[:n].cpu().numpy().astype(bool)
            res_h = result.cpu().numpy()
        if res_h[2] < 0:
            raise _lib.CytenB200Error("truncate_singular_values: the singular values contain non-finite entries")
        mults, secs = [], []
        for bi_large in range(leg.num_sectors):
            b0, b1 = leg.slices[bi_large]
            cnt = int(keep_h[b0:b1].sum())
            if cnt:
                mults.append(cnt)
                secs.append(leg.defining_sectors[bi_large])
        secs = np.asarray(secs, dtype=np.int64).reshape(-1, leg.symmetry.sector_ind_len)
        small_leg = ElementarySpace(leg.symmetry, secs, np.asarray(mults, dtype=np.int64), leg.is_dual)
        return keep_h, small_leg, float(res_h[0]), float(res_h[1])

    def apply_keep_mask(self, T: AbelianBackendData, axis: int, leg_col: int, leg: ElementarySpace,
                        keep: np.ndarray, small_leg: ElementarySpace) -> AbelianBackendData:
        """Projects a rank-2 (or rank-1 diagonal) block tensor onto the kept basis states of ``leg``
        (block_inds column ``leg_col``): _mask_contract / apply_mask_to_DiagonalTensor
        (abelian.cpp:2483-2582, 645-673).  Sectors without kept states are dropped and the leg's
        block indices are renumbered to ``small_leg``."""
        large_of_small = [i for i in range(leg.num_sectors) if keep[leg.slices[i, 0]:leg.slices[i, 1]].any()]
        renum = -np.ones(leg.num_sectors, dtype=np.int64)
        renum[large_of_small] = np.arange(len(large_of_small))
        sel = np.nonzero(renum[T.block_inds[:, leg_col]] >= 0)[0]
        diag = T.shapes.shape[1] == 1
        new_bi = T.block_inds[sel].copy()
        if diag:
            new_bi[:, :] = renum[new_bi[:, :1]]
        else:
            new_bi[:, leg_col] = renum[new_bi[:, leg_col]]
        idx_lists = [np.nonzero(keep[leg.slices[s, 0]:leg.slices[s, 1]])[0] for s in T.block_inds[sel, leg_col]]
        new_shapes = T.shapes[sel].copy()
        ax = 0 if diag else axis
        new_shapes[:, ax] = [len(ix) for ix in idx_lists]
        res, pos = self._sorted_data(T.dtype, new_bi, new_shapes)
        if len(sel) == 0:
            return res
        idx_ptr = np.concatenate([[0], np.cumsum([len(ix) for ix in idx_lists])])
        desc = np.zeros((len(sel), 6), dtype=np.int64)
        for q, i in enumerate(sel):
            m, n = (T.shapes[i, 0], 1) if diag else T.shapes[i]
            desc[q] = (m, n, T.offsets[i], res.offsets[pos[q]], idx_ptr[q], len(idx_lists[q]))
        with self._ctx():
            d_desc = to_device_i64(desc, self.device)
            d_idx = to_device_i64(np.concatenate(idx_lists) if idx_ptr[-1] else np.zeros(1, np.int64), self.device)
            self._call("cyb_gather_rows_cols", ptr(T.buf), ptr(res.buf), ptr(d_desc), ptr(d_idx), len(sel),
                       0 if diag else axis, T.cyb_dtype, self._stream())
        return res


class _Retyped:
    """A tensor view with replaced data (used for dtype promotion / re-laid-out operands)."""

    def __init__(self, t, data):
        self.data, self.codomain, self.domain = data, t.codomain, t.domain

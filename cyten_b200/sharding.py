"""Multi-GPU sharding of the abelian hot path: one process per GPU, charge sectors partitioned by
longest-processing-time-first (LPT) on flop cost.

The unit of distribution is the *block*: a block-sparse tensor that holds only a subset of its
symmetry-allowed blocks is still a valid tensor, and the full tensor is the SUM over ranks of the
rank-local subsets.  Every hot-path operation is (bi)linear in the blocks, so

* ``tdot`` / ``compose``: shard the operand along the sectors of a leg that stays OPEN -- every output
  block (``abelian.cpp:1424-1459``: independent per (row_a, col_b)) is then computed by exactly one rank,
  with no data-path collective;
* a chain of contractions (the DMRG ``H_eff`` matvec) stays rank-local for as long as the sharding leg
  stays open; when it is finally contracted the rank-local results are partial sums over that leg's
  sectors and ONE ``all_reduce`` on the common arena layout finishes the chain;
* ``svd`` / ``qr`` / ``eigh``: one independent factorisation per coupled sector
  (``abelian.cpp:3499-3541``), LPT by ``m*n*min(m,n)``; the only exchange step is the ``all_gather`` of the
  per-sector singular values (<= d*chi doubles) so that every rank evaluates the identical global cut
  (``tensor_backend.cpp:166-241``) and keeps the same per-sector dimensions.

``torch.distributed`` (NCCL over NVLink on the GPUs, gloo in the CPU tests) is only the plumbing.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import tensors as ct
from .backend import B200AbelianBackend, conventional_leg_order, _c_strides
from . import _lib
from .data import AbelianBackendData, ptr
from .spaces import TensorProduct
from .tensors import DiagonalTensor, SymmetricTensor, valid_block_inds


# ---- LPT ------------------------------------------------------------------------------------------
def lpt_assign(costs, nparts: int) -> np.ndarray:
    """Greedy longest-processing-time-first: items by descending cost (ties: lower index first) go to
    the currently least-loaded part (ties: lower part id).  Deterministic, so every rank computes the
    same assignment without communication.  Returns ``owner[len(costs)]``."""
    costs = np.asarray(costs, dtype=np.float64)
    order = np.lexsort((np.arange(len(costs)), -costs))
    load = np.zeros(nparts, dtype=np.float64)
    owner = np.zeros(len(costs), dtype=np.int64)
    for i in order:
        p = int(np.argmin(load))
        owner[i] = p
        load[p] += costs[i]
    return owner


def lpt_imbalance(costs, owner, nparts: int) -> float:
    """max load / mean load (1.0 = perfect)."""
    load = np.bincount(owner, weights=np.asarray(costs, dtype=np.float64), minlength=nparts)
    return float(load.max() / max(load.mean(), 1e-300))


class ProcessGroup:
    """Rank / world / group triple; ``group=None`` with world 1 is the single-GPU case."""

    def __init__(self, group=None):
        self.group = group
        if group is None and not (dist.is_available() and dist.is_initialized()):
            self.rank, self.world = 0, 1
        else:
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def all_reduce(self, t: torch.Tensor):
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_async(self, t: torch.Tensor):
        """Starts the all_reduce behind the work already queued on the current stream and returns a handle whose
        ``wait()`` orders the current stream after it; kernels launched in between run beside the collective."""
        if self.world > 1:
            return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return None

    def all_gather_into(self, out: torch.Tensor, t: torch.Tensor):
        if self.world > 1:
            dist.all_gather_into_tensor(out, t, group=self.group)
        else:
            out.copy_(t)
        return out


def attach_process_group(backend: B200AbelianBackend, group=None) -> ProcessGroup:
    pg = ProcessGroup(group)
    backend.shard = pg
    return pg


# ---- block subsets ---------------------------------------------------------------------------------
def take_blocks(T: SymmetricTensor, which) -> SymmetricTensor:
    """The tensor holding only blocks ``which`` (ascending positions) of T, in a new arena (one batched
    copy launch; the job list is a cached plan)."""
    be = T.backend
    which = np.asarray(which, dtype=np.int64)
    d = T.data
    key = ("take", d.sig, which.tobytes())
    plan = be._plan_get(key)
    if plan is None:
        res = be._empty(d.dtype, be.device, d.block_inds[which], d.shapes[which])
        prepared = (0, None)
        if len(which):
            sizes = d.sizes[which].reshape(-1, 1)
            ones = np.ones_like(sizes)
            prepared = be._prepare_copy(be._make_jobs(d.offsets[which], res.offsets, sizes, ones, ones))
        plan = dict(bi=res.block_inds, shapes=res.shapes, copy=prepared, sig=res.sig)
        if not be.dry_run:
            be._plan_put(key, plan)
    else:
        res = be._empty(d.dtype, be.device, plan["bi"], plan["shapes"])
        res._sig = plan["sig"]
    be._run_copy(d.buf, res.buf, plan["copy"], d.cyb_dtype)
    return SymmetricTensor(res, T.codomain, T.domain, be, T.labels)


def slice_leg(T: SymmetricTensor, leg, ranges: dict) -> SymmetricTensor:
    """T restricted along one leg to the index range ``ranges[sector] = (start, stop)`` of every listed sector
    (sectors not listed drop out): a tensor on a NARROWER leg (same sectors, smaller multiplicities).  Every
    operation is linear in each leg, so a tensor cut into such slabs along an open or a contracted leg (with the
    partner cut the same way) gives partial results that add up -- the finer-than-sector unit of the multi-GPU
    split.  One strided block-copy launch; the job list is a cached plan."""
    from .backend import _c_strides
    from .spaces import ElementarySpace, TensorProduct
    be = T.backend
    ax = T.get_leg_idcs([leg])[0]
    d = T.data
    key = ("slice", d.sig, ax, tuple(sorted((int(k), int(a), int(b)) for k, (a, b) in ranges.items())))
    plan = be._plan_get(key)
    ncod = T.num_codomain_legs
    old_leg = conventional_leg_order(T.codomain, T.domain)[ax]
    kept = np.asarray(sorted(int(k) for k in ranges), dtype=np.int64)
    if plan is None:
        remap = np.full(old_leg.num_sectors, -1, dtype=np.int64)
        remap[kept] = np.arange(len(kept))
        which = np.nonzero(remap[d.block_inds[:, ax]] >= 0)[0]
        bi = d.block_inds[which].copy()
        secs = bi[:, ax].copy()
        bi[:, ax] = remap[secs]
        start = np.asarray([ranges[int(q)][0] for q in secs], dtype=np.int64)
        stop = np.asarray([ranges[int(q)][1] for q in secs], dtype=np.int64)
        old_shapes = d.shapes[which]
        if np.any(start < 0) or np.any(stop > old_shapes[:, ax]) or np.any(stop <= start):
            raise ValueError("slice_leg: range outside the sector")
        shapes = old_shapes.copy()
        shapes[:, ax] = stop - start
        res = be._empty(d.dtype, be.device, bi, shapes)
        st_old = _c_strides(old_shapes)
        prepared = (0, None)
        if len(which):
            prepared = be._prepare_copy(be._make_jobs(d.offsets[which] + start * st_old[:, ax], res.offsets, shapes,
                                                      st_old, _c_strides(shapes)))
        plan = dict(bi=res.block_inds, shapes=res.shapes, copy=prepared, sig=res.sig)
        if not be.dry_run:
            be._plan_put(key, plan)
    else:
        res = be._empty(d.dtype, be.device, plan["bi"], plan["shapes"])
        res._sig = plan["sig"]
    be._run_copy(d.buf, res.buf, plan["copy"], d.cyb_dtype)
    new_leg = ElementarySpace(old_leg.symmetry, old_leg.defining_sectors[kept],
                              np.asarray([ranges[int(k)][1] - ranges[int(k)][0] for k in kept], dtype=np.int64), old_leg.is_dual)
    cod, dom = list(T.codomain.factors), list(T.domain.factors)
    if ax < ncod:
        cod[ax] = new_leg
    else:
        dom[T.num_legs - 1 - ax] = new_leg
    return SymmetricTensor(res, TensorProduct(cod, symmetry=T.symmetry), TensorProduct(dom, symmetry=T.symmetry), be, T.labels)


def column_ranges(costs, mults, world: int, heavy: float = 0.6, align: int = 16):
    """Work items for a split finer than sectors: a sector whose cost exceeds ``heavy`` x (one rank's share) is cut
    into equal index ranges (multiples of ``align``), the items go to ranks by LPT, and the parts of one sector
    that land on the same rank are made adjacent so that every rank holds at most ONE range per sector.
    Returns ``ranges[rank] = {sector: (start, stop)}``; identical on every rank (integer bookkeeping)."""
    costs = np.asarray(costs, dtype=np.float64)
    mults = np.asarray(mults, dtype=np.int64)
    ideal = float(costs.sum()) / world
    items = []                                           # (cost, sector, number of align-units)
    for s, (c, m) in enumerate(zip(costs, mults)):
        if c <= 0 or m <= 0:
            continue
        units = max(1, int(m) // align)
        nparts = int(min(units, np.ceil(c / (heavy * ideal)))) if c > heavy * ideal else 1
        base, extra = divmod(units, nparts)
        for k in range(nparts):
            u = base + (1 if k < extra else 0)
            items.append((c * u / units, s, u))
    owner = lpt_assign(np.asarray([it[0] for it in items]), world)
    ranges = [dict() for _ in range(world)]
    for s in sorted({it[1] for it in items}):
        m = int(mults[s])
        units_of_rank = np.zeros(world, dtype=np.int64)
        for (c, ss, u), o in zip(items, owner):
            if ss == s:
                units_of_rank[o] += u
        pos = 0
        holders = [r for r in range(world) if units_of_rank[r] > 0]
        for i, r in enumerate(holders):
            width = int(units_of_rank[r]) * align if i + 1 < len(holders) else m - pos      # the last holder takes the remainder
            ranges[r][s] = (pos, pos + width)
            pos += width
    return ranges


def sector_owner_of_leg(T: SymmetricTensor, leg, world: int):
    """LPT owner of every sector of leg ``leg`` (label or index) of T; cost of a sector = number of
    elements in the blocks carrying it (proportional to the flops those blocks cause downstream)."""
    ax = T.get_leg_idcs([leg])[0]
    nsec = conventional_leg_order(T.codomain, T.domain)[ax].num_sectors
    cost = np.bincount(T.data.block_inds[:, ax], weights=T.data.sizes.astype(np.float64), minlength=nsec)
    return ax, lpt_assign(cost, world), cost


def flop_costs_of_leg(T: SymmetricTensor, leg, chain) -> np.ndarray:
    """Flops that every sector of ``leg`` causes in ``chain(T_part)`` (a callable running the downstream
    contractions): a plan-only pass (``dry_run``: integer bookkeeping on the device, no arena is touched,
    nothing is cached) per sector, summing ``M*K*N`` over the matrix_dot lists of its composes.  This is
    the "flop-cost" of the LPT assignment; block volume alone underestimates the
    central sectors, whose blocks meet the largest partners."""
    be = T.backend
    ax = T.get_leg_idcs([leg])[0]
    nsec = conventional_leg_order(T.codomain, T.domain)[ax].num_sectors
    saved = (be.dry_run, be.plan_cache, be.plan_mkn)
    costs = np.zeros(nsec, dtype=np.float64)
    be.dry_run, be.plan_cache = True, False
    try:
        for s in range(nsec):
            mine = np.nonzero(T.data.block_inds[:, ax] == s)[0]
            if len(mine) == 0:
                continue
            be.plan_mkn = []
            chain(take_blocks(T, mine))
            costs[s] = float(sum(be.plan_mkn))
    finally:
        be.dry_run, be.plan_cache, be.plan_mkn = saved
    return costs


def flop_costs_total(T: SymmetricTensor, chain) -> float:
    """``sum M*K*N`` over the composes of ``chain(T)`` (plan-only pass, nothing cached, no arena touched)."""
    be = T.backend
    saved = (be.dry_run, be.plan_cache, be.plan_mkn)
    be.dry_run, be.plan_cache, be.plan_mkn = True, False, []
    try:
        chain(T)
        return float(sum(be.plan_mkn))
    finally:
        be.dry_run, be.plan_cache, be.plan_mkn = saved


def shard_by_leg(T: SymmetricTensor, leg, pg: ProcessGroup, costs=None):
    """Rank-local part of T: the blocks whose sector on ``leg`` is owned by this rank.
    Returns (local tensor, owner table of the leg's sectors)."""
    ax, owner, _ = sector_owner_of_leg(T, leg, pg.world)
    if costs is not None:
        owner = lpt_assign(np.asarray(costs, dtype=np.float64), pg.world)
    mine = np.nonzero(owner[T.data.block_inds[:, ax]] == pg.rank)[0]
    return take_blocks(T, mine), owner


def expand_to_full(T: SymmetricTensor) -> SymmetricTensor:
    """T re-laid-out on the arena of ALL symmetry-allowed blocks (valid_block_inds,
    abelian.cpp:329-399), absent blocks zero: the layout every rank shares."""
    be = T.backend
    # structure-only bookkeeping: cached like every other plan (one zero-fill + one copy launch per call)
    key = ("expand_full", T.data.sig, T.codomain.sig, T.domain.sig)
    plan = be._plan_get(key)
    if plan is not None:
        if plan.get("complete"):             # T already holds every allowed block: its arena IS the shared layout
            return T
        res = be._empty(T.data.dtype, be.device, plan["bi"], plan["shapes"], zero=True)
        res._sig = plan["sig"]
        be._run_copy(T.data.buf, res.buf, plan["copy"], T.data.cyb_dtype)
        return SymmetricTensor(res, T.codomain, T.domain, be, T.labels)
    full_bi = valid_block_inds(T.codomain, T.domain)
    legs = conventional_leg_order(T.codomain, T.domain)
    shapes = (np.stack([np.asarray(l.multiplicities)[full_bi[:, ax]] for ax, l in enumerate(legs)], axis=1)
              if legs else np.zeros((len(full_bi), 0), np.int64))
    d = T.data
    if d.num_blocks == len(full_bi) and np.array_equal(d.block_inds, full_bi):
        # Complete already (e.g. a column-split rank of the H_eff matvec computes a partial sum of EVERY output block): same
        # block list => same offsets (aligned_offsets is a function of the sizes), so no zero-fill and no copy -- 2 x 45 MB of
        # traffic per matvec at chi = 4096.  The padding between blocks is NOT zeroed in that case; nobody reads it.
        if not be.dry_run:
            be._plan_put(key, dict(complete=True))
        return T
    res = be._empty(T.data.dtype, be.device, full_bi, shapes, zero=True)
    if d.num_blocks:
        # position of every local block in the full (lexsorted) list: packed-key lookup
        nsec = np.asarray([l.num_sectors for l in legs], dtype=np.int64)
        strides = np.concatenate([[1], np.cumprod(nsec[:-1])]).astype(np.int64)   # last column primary
        key_full = full_bi @ strides
        pos = np.searchsorted(key_full, d.block_inds @ strides)
        if not np.array_equal(key_full[pos], d.block_inds @ strides):
            raise ValueError("expand_to_full: tensor has blocks that violate the charge rule")
        sizes = d.sizes.reshape(-1, 1)
        ones = np.ones_like(sizes)
        jobs = be._make_jobs(d.offsets, res.offsets[pos], sizes, ones, ones)
        prepared = be._prepare_copy(jobs)
        be._run_copy(d.buf, res.buf, prepared, d.cyb_dtype)
        be._plan_put(key, dict(bi=full_bi, shapes=shapes, copy=prepared, sig=res.sig))
    elif not be.dry_run:
        be._plan_put(key, dict(bi=full_bi, shapes=shapes, copy=(0, None), sig=res.sig))
    return SymmetricTensor(res, T.codomain, T.domain, be, T.labels)


def all_reduce_blocks(T: SymmetricTensor, pg: ProcessGroup) -> SymmetricTensor:
    """Sum of the rank-local tensors, replicated on every rank (one all_reduce of the arena)."""
    full = expand_to_full(T)
    if not T.backend.dry_run:
        pg.all_reduce(full.data.buf)
    return full


def all_gather_blocks(T: SymmetricTensor, pg: ProcessGroup) -> SymmetricTensor:
    """Replicates a tensor whose blocks are DISJOINTLY distributed over the ranks.  With disjoint
    supports the sum equals the union, and x + 0 is exact, so this is the all_reduce above."""
    return all_reduce_blocks(T, pg)


# ---- sharded operations ----------------------------------------------------------------------------
def tdot_sharded(A: SymmetricTensor, B: SymmetricTensor, legs1, legs2, pg: ProcessGroup, shard_leg=None):
    """``tdot`` with the OUTPUT blocks partitioned over the ranks: B is sharded along the sectors of one
    of its open legs (default: its last open leg), so each output block is computed by exactly one rank
    (no collective).  Returns the rank-local part of the result."""
    l2 = B.get_leg_idcs(legs2 if isinstance(legs2, (list, tuple)) else [legs2])
    if shard_leg is None:
        shard_leg = [i for i in range(B.num_legs) if i not in l2][-1]
    B_loc, _ = shard_by_leg(B, shard_leg, pg)
    return ct.tdot(A, B_loc, legs1, legs2)


class ShardedHEffective:
    """The two-site H_eff matvec (d_dmrg.py:70-88) on a theta sharded along ``vR``: the first three
    contractions leave ``vR`` open and are rank-local; the last one contracts ``vR`` with RP, so each
    rank holds a partial sum over its ``vR`` sectors and one all_reduce finishes the matvec."""

    def __init__(self, H, pg: ProcessGroup, shard_leg="vR", split_columns=False, column_align=16, column_heavy=0.6,
                 fused=False):
        self.H, self.pg, self.shard_leg = H, pg, shard_leg
        self.split_columns, self.column_align, self.column_heavy = split_columns, column_align, column_heavy
        self.fused = fused                  # rank-local chains run as GEMM / fused middle pass / GEMM (cyten_b200/fused.py)
        self.chains = [(self._op(H), None)]

    def _op(self, H):
        if not self.fused:
            return H
        from .fused import FusedHEffective
        return FusedHEffective(H)

    def shard_columns(self, theta: SymmetricTensor) -> SymmetricTensor:
        """Split finer than sectors: a heavy ``vR`` sector is cut into INDEX RANGES (``slice_leg``).  theta restricted
        to a range of vR indices is a tensor on a narrower leg; the first three contractions keep that leg open and
        the last contracts it with the same rows of RP, so rank r runs the unchanged chain on (theta_r, RP_r) and the
        partial results add up in the all_reduce.  No flops are duplicated and the GEMMs keep their fat dimensions
        (only the vR extent shrinks)."""
        from .workloads import HEffective
        H, pg = self.H, self.pg
        costs = flop_costs_of_leg(theta, self.shard_leg, H.matvec)
        self.costs = costs
        ax = theta.get_leg_idcs([self.shard_leg])[0]
        leg = conventional_leg_order(theta.codomain, theta.domain)[ax]
        self.ranges = column_ranges(costs, leg.multiplicities, pg.world, heavy=self.column_heavy, align=self.column_align)
        mine = self.ranges[pg.rank]
        self.slice_of = (self.shard_leg, mine)
        if not mine:                       # more ranks than work items: this rank contributes zeros
            self.chains = []
            self._empty_like = theta
            return take_blocks(theta, np.zeros(0, np.int64))
        theta_local = slice_leg(theta, self.shard_leg, mine)
        rp_ax = 0                          # RP[vR, wR | vR']: its first codomain leg is contracted with theta's vR
        RP_r = slice_leg(H.RP, rp_ax, mine)
        self.chains = [(self._op(HEffective(H.LP, RP_r, H.W1, H.W2, H.ct)), None)]
        return theta_local

    def shard(self, theta: SymmetricTensor) -> SymmetricTensor:
        """Rank-local part of theta.  Work items are the sectors of ``shard_leg`` weighted by the flops each causes
        in the matvec chain (plan-only passes); items go to ranks by LPT, identical on every rank (pure integer
        bookkeeping).  With ``split_columns`` the heavy sectors are cut into index ranges first (``shard_columns``)."""
        self.chains = [(self._op(self.H), None)]
        if self.pg.world == 1:
            return theta
        if self.split_columns:
            return self.shard_columns(theta)
        H, pg = self.H, self.pg
        costs = flop_costs_of_leg(theta, self.shard_leg, H.matvec)
        self.costs = costs
        ax = theta.get_leg_idcs([self.shard_leg])[0]
        items = [(float(c), sec) for sec, c in enumerate(costs) if c > 0]
        owner = lpt_assign(np.asarray([it[0] for it in items]), pg.world)
        self.items, self.item_owner = items, owner
        my_secs = sorted({it[1] for it, o in zip(items, owner) if o == pg.rank})
        return take_blocks(theta, np.nonzero(np.isin(theta.data.block_inds[:, ax], my_secs))[0])

    def matvec_local(self, theta_local: SymmetricTensor) -> SymmetricTensor:
        """This rank's partial sum of ``H_eff . theta`` on the shared all-blocks layout (sum over its chains)."""
        be = theta_local.backend
        acc = None
        for H_i, idx in self.chains:
            th = theta_local if idx is None else take_blocks(theta_local, idx)
            y = expand_to_full(H_i.matvec(th))
            if acc is None:
                acc = y
            elif not be.dry_run:
                n = acc.data.buf.numel() * (2 if acc.data.dtype == "complex128" else 1)
                with be._ctx():
                    be._call("cyb_axpby", ptr(y.data.buf), ptr(acc.data.buf), n, _lib.CYB_F64, 1.0, 1.0, be._stream())
        if acc is None:      # a rank without work (theta_local has no blocks): zeros on the full layout
            acc = expand_to_full(self.H.matvec(take_blocks(getattr(self, "_empty_like", theta_local), np.zeros(0, np.int64))))
        return acc

    def local_part(self, theta: SymmetricTensor) -> SymmetricTensor:
        """This rank's shard of a REPLICATED theta with the structure ``shard`` was called on (cached plans: one
        block-copy launch)."""
        if self.pg.world == 1:
            return theta
        if self.split_columns:
            mine = self.ranges[self.pg.rank]
            return slice_leg(theta, self.shard_leg, mine) if mine else take_blocks(theta, np.zeros(0, np.int64))
        ax = theta.get_leg_idcs([self.shard_leg])[0]
        my_secs = sorted({it[1] for it, o in zip(self.items, self.item_owner) if o == self.pg.rank})
        return take_blocks(theta, np.nonzero(np.isin(theta.data.block_inds[:, ax], my_secs))[0])

    def matvec_replicated(self, theta: SymmetricTensor) -> SymmetricTensor:
        """``H_eff . theta`` for a replicated theta, replicated result: the operator a Krylov solver sees
        (``cyten_b200.krylov.lanczos(SimpleNamespace(matvec=SH.matvec_replicated), theta)``).  Every rank runs the
        cheap vector operations on the full vectors redundantly (identical arithmetic, identical alpha/beta, no
        further collective); the matvec is the sharded chain + one all_reduce."""
        if self.pg.world == 1:
            return self.H.matvec(theta)
        return self.matvec(self.local_part(theta))

    def matvec(self, theta_local: SymmetricTensor) -> SymmetricTensor:
        """Replicated ``H_eff . theta`` from the rank-local shard of theta: one all_reduce of the result arena."""
        full = self.matvec_local(theta_local)
        if not full.backend.dry_run:
            self.pg.all_reduce(full.data.buf)
        return full


class OverlappedMatvec:
    """``ShardedHEffective.matvec`` with the all_reduce hidden behind the last GEMM.

    The result's ``vR`` leg comes from the open leg of RP.  Its sectors are cut into K contiguous groups; RP restricted
    to a group gives exactly the result blocks of that group, and because ``vR`` is the primary key of the block order
    those blocks are ONE contiguous range of the shared all-blocks arena.  A step is

        part 0:   x' = prefix(theta_r)   (LP . theta, fused middle pass -- once)      y_0 = x' . RP_0 -> range 0 of Y
        part g:                                                                         y_g = x' . RP_g -> range g of Y

    with ``all_reduce(Y[range g])`` started as soon as part g is queued: the collective of group g runs beside the GEMM
    of group g + 1 (NCCL's own stream), only the last group's all_reduce is exposed -- so the groups get decreasing
    shares of the arena (``fractions``).  No flop is duplicated (the groups partition the output blocks of the last
    GEMM) and the result equals ``ShardedHEffective.matvec`` up to the summation order of the all_reduce.
    ``Y`` is static (one per OverlappedMatvec): the parts are captured as one CUDA graph each (cyten_b200.graphs.
    CapturedPipeline) and the collectives are issued between the replays."""

    def __init__(self, SH: "ShardedHEffective", theta_local: SymmetricTensor, fractions=(0.6, 0.4)):
        if not SH.fused:
            raise ValueError("OverlappedMatvec needs the fused rank-local chain (ShardedHEffective(fused=True))")
        self.SH, self.pg = SH, SH.pg
        be = self.be = theta_local.backend
        self.chain = SH.chains[0][0] if SH.chains else None            # FusedHEffective over (LP, RP_r, W1, W2)
        # ---- the shared layout of the result: structure only (plan-only pass), identical on every rank ----
        was = be.dry_run
        be.dry_run = True
        try:
            if self.chain is not None:
                like = self.chain.matvec(theta_local)
            else:
                like = SH.H.matvec(take_blocks(getattr(SH, "_empty_like", theta_local), np.zeros(0, np.int64)))
        finally:
            be.dry_run = was
        self.codomain, self.domain, self.labels = like.codomain, like.domain, list(like.labels)
        legs = conventional_leg_order(like.codomain, like.domain)
        ax = like.get_leg_idcs(["vR"])[0]
        if ax != len(legs) - 1:
            raise ValueError("OverlappedMatvec: vR must be the last leg (the primary key of the block order)")
        full_bi = valid_block_inds(like.codomain, like.domain)
        shapes = np.stack([np.asarray(l.multiplicities)[full_bi[:, a]] for a, l in enumerate(legs)], axis=1)
        self.Y = be._empty(like.data.dtype, be.device, full_bi, shapes, zero=True)
        nsec = np.asarray([l.num_sectors for l in legs], dtype=np.int64)
        self._strides = np.concatenate([[1], np.cumprod(nsec[:-1])]).astype(np.int64)      # last column primary
        self._key_full = full_bi @ self._strides
        # ---- K contiguous groups of vR sectors holding about `fractions` of the arena ----
        sizes = np.prod(shapes, axis=1).astype(np.float64)
        per_sec = np.bincount(full_bi[:, ax], weights=sizes, minlength=int(nsec[ax]))
        cum = np.cumsum(per_sec) / max(per_sec.sum(), 1.0)
        cuts, acc = [0], 0.0
        for f in fractions[:-1]:
            acc += f
            c = int(np.searchsorted(cum, acc - 1e-12)) + 1
            cuts.append(min(max(c, cuts[-1] + 1), int(nsec[ax])))
        cuts.append(int(nsec[ax]))
        cuts = sorted(set(cuts))
        self.sector_groups = [(cuts[i], cuts[i + 1]) for i in range(len(cuts) - 1)]
        total = self.Y.buf.numel() if not be.dry_run else int(self.Y.offsets[-1] + sizes[-1]) if len(sizes) else 0
        self.block_ranges, self.ranges = [], []
        for s0, s1 in self.sector_groups:
            b0, b1 = int(np.searchsorted(full_bi[:, ax], s0)), int(np.searchsorted(full_bi[:, ax], s1))
            self.block_ranges.append((b0, b1))
            self.ranges.append((int(self.Y.offsets[b0]) if b0 < len(full_bi) else total,
                                int(self.Y.offsets[b1]) if b1 < len(full_bi) else total))
        # ---- RP restricted to each group (its open leg is its last leg) ----
        self.RP = []
        if self.chain is not None:
            RP = self.chain.H.RP
            col = RP.data.block_inds[:, RP.num_legs - 1]
            for s0, s1 in self.sector_groups:
                self.RP.append(take_blocks(RP, np.nonzero((col >= s0) & (col < s1))[0]))
        self._xp = None
        self.result = SymmetricTensor(self.Y, self.codomain, self.domain, be, self.labels)

    @property
    def num_parts(self):
        return len(self.sector_groups)

    def _expand_into(self, y: SymmetricTensor, g: int):
        be, Y, d = self.be, self.Y, y.data
        a, b = self.ranges[g]
        key = ("expand_into", d.sig, Y.sig, g, a, b)
        plan = be._plan_get(key)
        if plan is None:
            prepared = (0, None)
            if d.num_blocks:
                keys = d.block_inds @ self._strides
                pos = np.searchsorted(self._key_full, keys)
                if not np.array_equal(self._key_full[pos], keys):
                    raise ValueError("OverlappedMatvec: a part has blocks that violate the charge rule")
                b0, b1 = self.block_ranges[g]
                if pos.min() < b0 or pos.max() >= b1:
                    raise ValueError("OverlappedMatvec: a part wrote outside its range of the arena")
                sizes = d.sizes.reshape(-1, 1)
                ones = np.ones_like(sizes)
                prepared = be._prepare_copy(be._make_jobs(d.offsets, Y.offsets[pos], sizes, ones, ones))
            plan = dict(copy=prepared, complete=d.num_blocks == self.block_ranges[g][1] - self.block_ranges[g][0])
            if not be.dry_run:
                be._plan_put(key, plan)
        if be.dry_run:
            return
        if not plan["complete"] and b > a:          # the previous step's all_reduce left sums in the absent blocks
            Y.buf[a:b].zero_()
        be._run_copy(d.buf, Y.buf, plan["copy"], d.cyb_dtype)

    def part(self, g: int, theta_local: SymmetricTensor):
        """Queues part g of a step (kernel launches only once the plans are cached: capturable)."""
        a, b = self.ranges[g]
        if self.chain is None:                      # a rank without work contributes zeros
            if not self.be.dry_run and b > a:
                self.Y.buf[a:b].zero_()
            return
        if g == 0:
            self._xp = self.chain.prefix(theta_local)
        self._expand_into(self.chain.suffix(self._xp, self.RP[g]), g)

    def reduce_async(self, g: int):
        a, b = self.ranges[g]
        return self.pg.all_reduce_async(self.Y.buf[a:b]) if b > a and not self.be.dry_run else None

    def __call__(self, theta_local: SymmetricTensor, run_part=None) -> SymmetricTensor:
        """One step; ``run_part(g)`` replaces the eager part g (a graph replay)."""
        works = []
        for g in range(self.num_parts):
            if run_part is None:
                self.part(g, theta_local)
            else:
                run_part(g)
            works.append(self.reduce_async(g))
        for w in works:
            if w is not None:
                w.wait()
        return self.result


def svd_costs(tens: SymmetricTensor) -> np.ndarray:
    s = tens.data.shapes
    return (s[:, 0] * s[:, 1] * np.minimum(s[:, 0], s[:, 1])).astype(np.float64)


def gather_singular_values(be: B200AbelianBackend, s: AbelianBackendData, leg, owner_of_sector, pg: ProcessGroup):
    """NCCL ``all_gather`` of the per-sector singular values (section 8(e)): every rank contributes its
    sectors' values as one padded segment; the result is the dense S vector over ``leg`` (internal basis
    order, diagonal_tensor_to_block, abelian.cpp:1678-1692), identical on every rank."""
    mults = np.asarray(leg.multiplicities, dtype=np.int64)
    nsec = leg.num_sectors
    per_rank = [np.nonzero(owner_of_sector == r)[0] for r in range(pg.world)]
    seg_len = [int(mults[idx].sum()) for idx in per_rank]
    pad = max(max(seg_len), 1)
    send = torch.zeros(pad, dtype=torch.float64, device=be.device)
    # pack my blocks in ascending sector order
    mine = per_rank[pg.rank]
    start_in_seg = {int(sec): int(o) for sec, o in zip(mine, np.concatenate([[0], np.cumsum(mults[mine])])[:-1])}
    if s.num_blocks:
        sizes = s.shapes[:, :1]
        dst = np.asarray([start_in_seg[int(sec)] for sec in s.block_inds[:, 0]], dtype=np.int64)
        ones = np.ones_like(sizes)
        be._copy_blocks(s.buf, send, be._make_jobs(s.offsets, dst, sizes, ones, ones), 0)
    recv = torch.empty(pad * pg.world, dtype=torch.float64, device=be.device)
    pg.all_gather_into(recv, send)
    dense = torch.zeros(max(leg.dim, 1), dtype=torch.float64, device=be.device)
    src, dst, sz = [], [], []
    for r, idx in enumerate(per_rank):
        o = 0
        for sec in idx:
            src.append(r * pad + o)
            dst.append(int(leg.slices[sec, 0]))
            sz.append(int(mults[sec]))
            o += int(mults[sec])
    if sz:
        sz = np.asarray(sz, dtype=np.int64).reshape(-1, 1)
        ones = np.ones_like(sz)
        be._copy_blocks(recv, dense, be._make_jobs(np.asarray(src), np.asarray(dst), sz, ones, ones), 0)
    return dense


def truncated_svd_sharded(T: SymmetricTensor, pg: ProcessGroup, new_labels=("a", "b", "c", "d"),
                          new_leg_dual=False, normalize_to=None, chi_max=None, chi_min=1, degeneracy_tol=0.0,
                          trunc_cut=0.0, svd_min=0.0):
    """``truncated_svd`` (decompositions.cpp:672-709) with the coupled sectors LPT-partitioned over the
    ranks.  T is replicated; each rank factorises its sectors, the singular values are all-gathered, the
    global cut is evaluated identically everywhere, and each rank projects ITS blocks of U, S, Vh.
    Returns rank-local (U, S, Vh) (disjoint block sets; ``all_gather_blocks`` replicates them), err,
    renormalize, and the sector owner table."""
    be: B200AbelianBackend = T.backend
    tens, new_co_domain, cc, cd = ct._decomposition_prepare(T, new_leg_dual)
    leg = new_co_domain.factors[0]
    rows = be._matched_sectors(tens, leg)
    # cost per new-leg sector; sectors without a theta block cost nothing and go to rank 0's tail
    cost = np.zeros(leg.num_sectors, dtype=np.float64)
    blk_cost = svd_costs(tens)
    for (j, n, k, i) in rows:
        if i >= 0:
            cost[n] = blk_cost[i]
    owner = lpt_assign(cost, pg.world)
    owned_rows = [t for t, r in enumerate(rows) if owner[r[1]] == pg.rank]
    u, s, vh = be.svd(tens, new_co_domain, only_rows=owned_rows)
    dense = gather_singular_values(be, s, leg, owner, pg)
    keep, small_leg, err, new_norm, S_norm = be.truncate_dense(dense, leg, chi_max=chi_max, chi_min=chi_min,
                                                               degeneracy_tol=degeneracy_tol, trunc_cut=trunc_cut,
                                                               svd_min=svd_min)
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
        U = ct.split_legs(U, [0])
    if cd:
        Vh = ct.split_legs(Vh, [1])
    return U, S, Vh, err, renormalize, owner


# ---- BASELINE config 5 on N GPUs -------------------------------------------------------------------------------------
def all_gather_diagonal(S: DiagonalTensor, pg: ProcessGroup) -> DiagonalTensor:
    """Replicates a DiagonalTensor whose blocks are disjointly distributed over the ranks: every sector of the leg gets a
    block (zeros where no rank holds one), the local blocks are written at their places, one all_reduce."""
    be, leg = S.backend, S.leg
    ns = leg.num_sectors
    bi = np.stack([np.arange(ns), np.arange(ns)], axis=1).astype(np.int64)
    data = AbelianBackendData.empty("float64", be.device, bi, np.asarray(leg.multiplicities, np.int64).reshape(-1, 1), zero=True)
    for n, row in enumerate(S.data.block_inds):
        data.block(int(row[0])).copy_(S.data.block(n))
    pg.all_reduce(data.buf)
    return DiagonalTensor(data, leg, be, S.labels)


class ShardedSweepOps:
    """``SweepStep`` operations on N ranks: the truncated SVD runs sector-sharded (global cut from the all-gathered
    singular values), its factors are replicated afterwards; the environment updates are sharded along the leg that stays
    open through their three contractions (disjoint result blocks, one gather each); the scale_axis calls run replicated."""

    def __init__(self, pg: ProcessGroup):
        from .workloads import NativeOps
        self.pg = pg
        self._n = NativeOps()

    def __getattr__(self, name):
        return getattr(self._n, name)

    def shard(self, T, leg):
        """Block subset of T owned by this rank along an OPEN leg (LPT by block volume): the environment-update chains
        keep that leg open to the end, so the ranks produce disjoint blocks of the result."""
        return shard_by_leg(T, leg, self.pg)[0]

    def gather(self, T):
        return all_gather_blocks(T, self.pg)

    def truncated_svd(self, theta, new_labels, chi_max, svd_min):
        a, b = new_labels
        U, S, Vh, err, _, _ = truncated_svd_sharded(theta, self.pg, new_labels=(a, a, b, b), chi_max=chi_max, svd_min=svd_min)
        return all_gather_blocks(U, self.pg), all_gather_diagonal(S, self.pg), all_gather_blocks(Vh, self.pg), err


def shard_sweep_step(step, pg: ProcessGroup, split_columns=True, column_heavy=0.35, column_align=16):
    """Turns a ``cyten_b200.workloads.SweepStep`` into its N-rank version: H_eff matvec sharded along ``vR`` (replicated
    in / out, one all_reduce), sharded truncated SVD, replicated environment updates."""
    if pg.world == 1:
        return step
    from types import SimpleNamespace
    step.ops = ShardedSweepOps(pg)
    plain_heff = step.heff
    cache = {}

    def heff():
        if "SH" not in cache:
            fused, step.fused = bool(getattr(step, "fused", False)), False       # the rank-local chains are fused instead
            try:
                H = plain_heff()
            finally:
                step.fused = fused
            cache["SH"] = ShardedHEffective(H, pg, split_columns=split_columns, column_heavy=column_heavy,
                                           column_align=column_align, fused=fused)

        def matvec(theta):
            SH = cache["SH"]
            if "ready" not in cache:
                SH.shard(theta)              # work partition for this theta structure (plan-only passes, once)
                cache["ready"] = True
            return SH.matvec_replicated(theta)
        return SimpleNamespace(matvec=matvec)
    step.heff = heff
    return step

"""Fused two-site effective Hamiltonian: GEMM, ONE streaming pass, GEMM.

The reference's ``HEffective.matvec`` (toycodes/tenpy_toycodes/d_dmrg.py:70-88) -- and the call-for-call mirror
``cyten_b200.workloads.HEffective`` -- is 5 ``permute_legs`` + 4 ``compose``: ``LP . theta``, two MPO applications, ``. RP``.
The MPO applications contract legs of dimension D d = 10 but move the five-times-theta-sized intermediate through HBM five
times (compose + permute, twice, plus the re-layout for the last GEMM).  ``PlanarHEffective`` (d_dmrg.py:94-116,
src/tensors/planar.cpp) hands the whole diagram to a planner; this is the device work-list such a planner emits for it:

    x  = LP . theta                       grouped DMMA GEMM       -> x [vL, (wL p0 p1), vR]
    x' = (W1 . W2) acting on (wL p0 p1)   cyb_middle_apply        -> x'[vL, (p0 p1 wR), vR]     one read + one write
    y  = x' . RP                          grouped DMMA GEMM

For a fixed pair of (vL, vR) sectors the blocks of x are the rows of a small dense operator -- W1 . W2 contracted over wC,
restricted to the middle-leg charges that conserve the total -- which is the same for every (vL index, vR index); the
operator tables are built once per structure (integer bookkeeping from ``block_inds`` + a 20 x 20 einsum on the host).
The result equals the unfused chain up to the summation order of the <= 8 terms per element.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from . import tensors as ct
from .backend import conventional_leg_order
from .data import AbelianBackendData, ptr, to_device_i64
from .tensors import SymmetricTensor


class FusedHEffective:
    """``matvec`` of a ``cyten_b200.workloads.HEffective`` (bent LP / RP / W1 / W2) with the middle fused."""

    def __init__(self, H):
        self.H = H
        self.be = H.LP.backend
        self._plans = {}
        # dense MPO tensors on the host (tiny): axes = labels [p', wOut, p, wIn]
        self._w1 = H.W1.to_dense()
        self._w2 = H.W2.to_dense()
        if list(H.W1.labels) != ["p0", "wC", "p0*", "wL"] or list(H.W2.labels) != ["p1", "wR", "p1*", "wC"]:
            raise ValueError("FusedHEffective expects the bent MPO tensors of HEffective (labels p0 wC p0* wL / p1 wR p1* wC)")
        # operator on the middle legs: Wc[p0', p1', wR ; wL, p0, p1]
        self._wc = np.einsum("acdb,efgc->aefbdg", self._w1, self._w2)

    # ---- the unfused middle (plan-only): gives the structure of x' exactly as the second GEMM wants it ----
    def _middle_unfused(self, x):
        H = self.H
        x = ct.permute_legs(x, ["wL*", "p0"], ["vL", "vR", "p1"])
        x = ct.compose(H.W1, x)
        x = ct.permute_legs(x, ["wC", "p1"], ["p0", "vL", "vR"])
        x = ct.compose(H.W2, x)
        return ct.permute_legs(x, ["vL", "p0", "p1"], ["vR", "wR"])

    def _plan(self, x: SymmetricTensor):
        be = self.be
        key = x.data.sig
        plan = self._plans.get(key)
        if plan is not None:
            return plan
        if list(x.labels) != ["vL", "wL*", "p0", "p1", "vR"] or x.num_codomain_legs != 2:
            raise ValueError(f"FusedHEffective: unexpected intermediate {x.labels}")
        # structure of the result from a plan-only pass of the unfused sequence
        was = be.dry_run
        be.dry_run = True
        try:
            y_like = self._middle_unfused(x)
            y_bi, y_shapes = y_like.data.block_inds.copy(), y_like.data.shapes.copy()
            y_cod, y_dom, y_labels = y_like.codomain, y_like.domain, list(y_like.labels)
        finally:
            be.dry_run = was
        if y_labels != ["vL", "p0", "p1", "wR", "vR"]:
            raise ValueError(f"FusedHEffective: unexpected result legs {y_labels}")
        Y = be._empty(x.data.dtype, be.device, y_bi, y_shapes)       # offsets of the result arena
        xl = conventional_leg_order(x.codomain, x.domain)        # (vL, wL, p0, p1, vR)
        yl = conventional_leg_order(y_cod, y_dom)                # (vL, p0', p1', wR, vR)
        X = x.data
        lib = be._library if be._library is not None else _lib.load()
        max_rows, ucols = int(lib.cyb_middle_apply_max_rows()), int(lib.cyb_middle_apply_unit_cols())
        cx = X.dtype == "complex128"
        wc = self._wc if cx or not np.iscomplexobj(self._wc) else self._wc.astype(np.complex128)
        groups_in, groups_out = {}, {}
        for n, row in enumerate(X.block_inds):
            groups_in.setdefault((int(row[0]), int(row[4])), []).append(n)
        for n, row in enumerate(y_bi):
            groups_out.setdefault((int(row[0]), int(row[4])), []).append(n)
        G, D, Cf, W0 = [], [], [], [0]
        for (sa, sb), outs in sorted(groups_out.items()):
            ins = groups_in.get((sa, sb), [])
            na, nb = int(xl[0].multiplicities[sa]), int(xl[4].multiplicities[sb])
            in_rows, in_idx = [], []
            for n in ins:
                _, sw, s0, s1, _ = (int(v) for v in X.block_inds[n])
                mw, m0, m1 = (int(v) for v in X.shapes[n, 1:4])
                kblk = mw * m0 * m1
                for iw in range(mw):
                    for i0 in range(m0):
                        for i1 in range(m1):
                            kloc = (iw * m0 + i0) * m1 + i1
                            in_rows.append((int(X.offsets[n]) + kloc * nb, kblk * nb))
                            in_idx.append((int(xl[1].slices[sw, 0]) + iw, int(xl[2].slices[s0, 0]) + i0, int(xl[3].slices[s1, 0]) + i1))
            out_rows, out_idx = [], []
            for n in outs:
                _, t0, t1, tw, _ = (int(v) for v in y_bi[n])
                m0, m1, mw = (int(v) for v in y_shapes[n, 1:4])
                mblk = m0 * m1 * mw
                for j0 in range(m0):
                    for j1 in range(m1):
                        for jw in range(mw):
                            mloc = (j0 * m1 + j1) * mw + jw
                            out_rows.append((int(Y.offsets[n]) + mloc * nb, mblk * nb))
                            out_idx.append((int(yl[1].slices[t0, 0]) + j0, int(yl[2].slices[t1, 0]) + j1, int(yl[3].slices[tw, 0]) + jw))
            if len(in_rows) > max_rows or len(out_rows) > max_rows:
                raise _lib.CytenB200Error(f"FusedHEffective: operator of {len(out_rows)} x {len(in_rows)} rows per sector pair "
                                          f"exceeds cyb_middle_apply_max_rows() = {max_rows}")
            coef = np.zeros((len(out_rows), len(in_rows)), dtype=np.complex128 if cx else np.float64)
            for o, (a0, a1, a2) in enumerate(out_idx):
                for i, (b0, b1, b2) in enumerate(in_idx):
                    coef[o, i] = wc[a0, a1, a2, b0, b1, b2]
            G.append((na, nb, len(in_rows), len(out_rows), len(D), sum(c.size for c in Cf)))
            D.extend(in_rows + out_rows)
            Cf.append(coef.reshape(-1))
            W0.append(W0[-1] + na * ((nb + ucols - 1) // ucols))
        plan = dict(y_bi=y_bi, y_shapes=y_shapes, y_cod=y_cod, y_dom=y_dom, y_labels=y_labels, ngroups=len(G), nwork=W0[-1],
                    max_rows=max([max(g[2], g[3]) for g in G], default=0),
                    bytes=sum((g[2] + g[3]) * g[0] * g[1] for g in G) * (16 if cx else 8), sig=None)
        if G and not be.dry_run:
            with be._ctx():
                plan["groups"] = to_device_i64(np.asarray(G, dtype=np.int64), be.device)
                plan["desc"] = to_device_i64(np.asarray(D, dtype=np.int64), be.device)
                coef_all = np.concatenate(Cf) if Cf else np.zeros(1)
                plan["coef"] = torch.from_numpy(np.ascontiguousarray(coef_all)).to(be.device)
                plan["work0"] = to_device_i64(np.asarray(W0, dtype=np.int64), be.device)
        if not be.dry_run:
            self._plans[key] = plan
        return plan

    def middle(self, x: SymmetricTensor) -> SymmetricTensor:
        """x[vL, wL* | p0, p1, vR] -> x'[vL, p0, p1 | vR, wR]: both MPO tensors applied in one pass."""
        be = self.be
        plan = self._plan(x)
        Y = be._empty(x.data.dtype, be.device, plan["y_bi"], plan["y_shapes"])
        Y._sig = plan["sig"]
        if plan["sig"] is None:
            plan["sig"] = Y.sig
        if plan["ngroups"] and not be.dry_run:
            with be._ctx():
                be._call("cyb_middle_apply", ptr(x.data.buf), ptr(Y.buf), ptr(plan["groups"]), ptr(plan["desc"]), ptr(plan["coef"]),
                         ptr(plan["work0"]), plan["ngroups"], plan["nwork"], plan["max_rows"], Y.cyb_dtype, be._stream())
        return SymmetricTensor(Y, plan["y_cod"], plan["y_dom"], be, plan["y_labels"])

    def prefix(self, theta: SymmetricTensor) -> SymmetricTensor:
        """Everything before ``vR`` is contracted: ``LP . theta`` and the fused middle pass -> x'[vL, p0, p1 | vR, wR]."""
        x = ct.permute_legs(theta, ["vL"], ["vR", "p1", "p0"])
        x = ct.compose(self.H.LP, x)                                # vL wL* p0 p1 vR
        return self.middle(x)                                       # vL p0 p1 wR vR

    def suffix(self, x: SymmetricTensor, RP=None) -> SymmetricTensor:
        """``x' . RP`` (``RP``: the right environment or a subset of its blocks -- the multi-GPU path runs this GEMM in
        groups of output sectors so that each group's all_reduce hides behind the next group's GEMM)."""
        x = ct.compose(x, self.H.RP if RP is None else RP)          # vL p0 p1 vR
        return ct.permute_legs(x, domain=["vR", "p1"])

    def matvec(self, theta: SymmetricTensor) -> SymmetricTensor:
        return self.suffix(self.prefix(theta))

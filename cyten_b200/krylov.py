"""Lanczos ground-state search over the device vector operations (the loop the hot path
sits in).  Mirrors ``cyten.krylov_based.LanczosGroundState`` (src/tensors/krylov_based.cpp:800-960: options
``N_min``, ``N_max``, ``N_cache``, ``P_tol``, ``E_tol``, ``min_gap``, ``cutoff``, ``reortho``, ``E_shift``; same
recurrence, same Ritz-residual convergence test, same rebuild of the result when fewer than N vectors are cached).

B200-first differences: an iteration is  v = u/b | w = H v | alpha = Re<v, w> | u' = w - alpha v - b v_prev, ||u'||^2
where alpha never leaves the device (the three-term recurrence and the norm of the new vector are ONE fused kernel
reading alpha from device memory), so there is exactly one host synchronisation per iteration: the 16-byte read of
(alpha, ||u'||^2) that the host-side convergence test (eigh of the k x k tridiagonal matrix) needs.  ``GraphedOperator``
replays the plan-cached matvec as one CUDA graph.
"""
from __future__ import annotations

import numpy as np
import torch

from .graphs import CapturedStep
from .tensors import SymmetricTensor, linear_combination, norm, scalar_multiply


class TensorLinearOperator:
    """``matvec(v) = tdot(H, v, which_leg, 0)`` (cyten.sparse.TensorLinearOperator)."""

    def __init__(self, tensor: SymmetricTensor, which_leg=-1):
        from . import tensors as ct
        self.tensor, self.which_leg, self._ct = tensor, which_leg, ct

    def matvec(self, vec: SymmetricTensor) -> SymmetricTensor:
        return self._ct.tdot(self.tensor, vec, [self.which_leg], [0])


class GraphedOperator:
    """Wraps an operator whose ``matvec`` keeps the structure of its argument: the first call with a given structure
    captures the (plan-cached) launch sequence into a CUDA graph, later calls copy the argument into the graph's
    static input arena and replay.  The returned tensor is the graph's static output: it is overwritten by the next
    ``matvec`` (the Lanczos recurrence consumes it before that)."""

    def __init__(self, op, warmup: int = 3):
        self.op, self.warmup, self._steps = op, warmup, {}

    def _step(self, vec: SymmetricTensor) -> CapturedStep:
        step = self._steps.get(vec.data.sig)
        if step is None:
            static = SymmetricTensor(_clone(vec.data), vec.codomain, vec.domain, vec.backend, vec.labels)
            step = self._steps[vec.data.sig] = CapturedStep(self.op.matvec, static, warmup=self.warmup)
        return step

    def input_like(self, vec: SymmetricTensor) -> SymmetricTensor:
        """The graph's static input tensor for vectors of this structure: a caller that PRODUCES the vector (the
        normalisation of a Lanczos step) writes it there directly and saves the copy."""
        return self._step(vec).input

    def matvec(self, vec: SymmetricTensor) -> SymmetricTensor:
        step = self._step(vec)
        if vec is not step.input:
            step.input.data.buf.copy_(vec.data.buf)
        return step()


def _clone(data):
    from .data import AbelianBackendData
    return AbelianBackendData(data.dtype, data.device, data.buf.clone(), data.block_inds, data.offsets, data.shapes,
                              sig=data._sig)


class LanczosGroundState:
    """``E0, psi0, N = LanczosGroundState(H, psi_init, options).run()``."""

    def __init__(self, H, psi0: SymmetricTensor, options: dict | None = None):
        o = dict(options or {})
        self.H, self.psi0 = H, psi0
        self.N_min = int(o.pop("N_min", 2))
        self.N_max = int(o.pop("N_max", 20))
        self.N_cache = int(o.pop("N_cache", self.N_max))
        self.P_tol = float(o.pop("P_tol", 1.0e-14))
        self.E_tol = float(o.pop("E_tol", np.inf))
        self.min_gap = float(o.pop("min_gap", 1.0e-12))
        self.cutoff = float(o.pop("cutoff", np.finfo(np.float64).eps * 100))
        self.reortho = bool(o.pop("reortho", False))
        self.E_shift = o.pop("E_shift", None)
        if o:
            raise ValueError(f"unknown Lanczos options: {sorted(o)}")
        if self.N_min < 2:
            raise ValueError("Should perform at least 2 steps.")
        if self.N_cache < 2:
            raise ValueError("Need to cache at least two vectors.")
        self.be = psi0.backend
        n = self.N_max + 1
        self.h_krylov = np.zeros((n, n))
        self.Es = np.zeros((self.N_max, self.N_max))
        self._result_krylov = np.ones(1)
        self._cache: list = []                                   # (u_k, beta_k): v_k = u_k / beta_k
        self.host_syncs = 0
        self._psi0_norm = None
        dev = self.be.device
        self._scal = torch.zeros(4, dtype=torch.float64, device=dev)      # (alpha, 0 | ||u'||^2, 0)
        self._ov = torch.zeros(2, dtype=torch.float64, device=dev)

    # -- pieces ---------------------------------------------------------------------------------------
    def _matvec(self, v):
        w = self.H.matvec(v)
        if self.E_shift is not None:
            w = linear_combination(1.0, w, float(self.E_shift), v)
        return w

    def _to_cache(self, v):
        self._cache.append(v)
        if len(self._cache) > self.N_cache:
            self._cache.pop(0)

    def _wrap(self, data, like):
        return SymmetricTensor(data, like.codomain, like.domain, like.backend, like.labels)

    def _same_structure(self, w, v):
        """The recurrence runs on arenas of identical layout; a matvec result with another block set (blocks H
        cannot reach, or new ones) is brought to v's layout + its own extra blocks by one linear_combination."""
        return w if w.data.sig == v.data.sig else linear_combination(1.0, w, 0.0, v)

    def _normalised(self, u, b):
        """v = u / b, written straight into the operator's static input arena when it has one (GraphedOperator)."""
        static = getattr(self.H, "input_like", None)
        if static is not None and self.E_shift is None:
            v = static(u)
            self.be.lincomb3_dev([u.data, None, None], [1.0 / b, 0.0, 0.0], out=v.data)
            return v
        return u if b == 1.0 else scalar_multiply(1.0 / b, u)

    def _three_term(self, w, v, prev, b):
        """alpha = Re<v, w> (device) ; u' = w - alpha v - b v_prev ; ||u'||^2 -- two launches, no sync.
        ``prev = (u_prev, b_prev)``: the cache holds the UNNORMALISED vectors, v_prev = u_prev / b_prev."""
        be = self.be
        alpha, ss = self._scal[0:2], self._scal[2:3]
        be.inner_dev(v.data, w.data, True, real_only=True, out=alpha)
        same = w.data.sig == v.data.sig and (prev is None or prev[0].data.sig == v.data.sig)
        if same:
            ops = [w.data, v.data, None if prev is None else prev[0].data]
            u = be.lincomb3_dev(ops, [1.0, -1.0, 0.0 if prev is None else -b / prev[1]], coef_dev=[None, alpha, None], sumsq=ss)
            return self._wrap(u, v)
        # general block structures (w gained or lost blocks): reference semantics through linear_combination
        a = float(alpha[0].item())
        self.host_syncs += 1
        u = linear_combination(1.0, w, -a, v)
        if prev is not None:
            u = linear_combination(1.0, u, -b / prev[1], prev[0])
        be.norm_sq_dev(u.data, out=self._scal[2:4])
        return u

    def _reorthogonalise(self, u):
        """u -= <v_i, u> v_i for the cached vectors but the last: <v_i, u> v_i = <u_i, u> u_i / b_i^2, the raw overlap
        stays on the device."""
        be = self.be
        for c, bc in self._cache[:-1]:
            if c.data.sig == u.data.sig:
                be.inner_dev(c.data, u.data, True, out=self._ov)
                be.lincomb3_dev([u.data, c.data, None], [1.0, -1.0 / bc ** 2, 0.0], coef_dev=[None, self._ov, None], out=u.data)
            else:
                u = linear_combination(1.0, u, -be.inner(c.data, u.data) / bc ** 2, c)
        if self._cache[:-1]:
            be.norm_sq_dev(u.data, out=self._scal[2:4])
        return u

    def _calc_result_krylov(self, k):
        if k == 0:
            self.Es[0, 0] = self.h_krylov[0, 0]
            self._result_krylov = np.ones(1)
            return
        E, V = np.linalg.eigh(self.h_krylov[:k + 1, :k + 1])
        self.Es[k, :k + 1] = E
        self._result_krylov = V[:, 0].copy()

    def _converged(self, k) -> bool:
        ritz = abs(self._result_krylov[k]) * abs(self.h_krylov[k, k + 1])
        gap = max(self.Es[k, 1] - self.Es[k, 0], self.min_gap)
        return (ritz / gap) ** 2 < self.P_tol and (self.Es[k - 1, 0] - self.Es[k, 0]) < self.E_tol

    # -- the iteration ----------------------------------------------------------------------------------
    def _build_krylov(self) -> int:
        u = self.psi0
        beta = norm(u)
        self.host_syncs += 1
        if beta < self.cutoff:
            raise ValueError(f"Norm of self.psi0 too small: {beta}")
        self.psi0 = u = scalar_multiply(1.0 / beta, u)
        if self._psi0_norm is None:
            self._psi0_norm = beta                                  # only needed by LanczosEvolution
        beta, performed = 1.0, 0
        v_next = None
        for k in range(self.N_max):
            v = v_next if v_next is not None else self._normalised(u, beta)
            self._to_cache((u, beta))                              # v_k = u_k / beta_k
            w = self._same_structure(self._matvec(v), v)
            prev = self._cache[-2] if (k > 0 and not self.reortho) else None
            u = self._three_term(w, v, prev, beta)
            if self.reortho:
                u = self._reorthogonalise(u)
            alpha, _, ss, _ = self._scal.tolist()                  # the one synchronisation of the iteration
            self.host_syncs += 1
            beta = float(np.sqrt(max(ss, 0.0)))
            # the next normalisation is enqueued BEFORE the host-side convergence test (eigh of the tridiagonal matrix),
            # so the GPU is not idle during it; if the test then stops the loop the kernel was wasted, nothing else
            v_next = self._normalised(u, beta) if (abs(beta) >= self.cutoff and k + 1 < self.N_max) else None
            self.h_krylov[k, k] = alpha
            self._calc_result_krylov(k)
            self.h_krylov[k, k + 1] = self.h_krylov[k + 1, k] = beta
            performed = k + 1
            if abs(beta) < self.cutoff or (k + 1 >= self.N_min and self._converged(k)):
                break
        return performed

    def _rebuild(self, psif, n_rebuild):
        """Vectors that fell out of the cache are regenerated with the known alpha / beta: no synchronisation."""
        vf = self._result_krylov
        v, v_prev, beta, vs = self.psi0, None, 0.0, []
        for k in range(n_rebuild):
            w = self._same_structure(self._matvec(v), v)
            u = linear_combination(1.0, w, -self.h_krylov[k, k], v)
            if self.reortho:
                # the reference's rebuild pushes v_k to the cache first and loops over all but the last entry
                # (krylov_based.cpp:902-910): N_cache - 1 vectors, the same set the build phase orthogonalised against
                for c in (vs[-(self.N_cache - 1):] if self.N_cache > 1 else []):
                    u = linear_combination(1.0, u, -self.be.inner(c.data, u.data), c)
            elif k > 0:
                u = linear_combination(1.0, u, -beta, v_prev)
            vs.append(v)
            beta = self.h_krylov[k, k + 1]
            v_prev, v = v, scalar_multiply(1.0 / beta, u)
            psif = linear_combination(1.0, psif, vf[k + 1], v)
        return psif

    def _calc_result_full(self, N):
        vf = self._result_krylov
        psif = scalar_multiply(vf[0], self.psi0)
        len_cache = len(self._cache)
        for k in range(1, min(len_cache + 1, N)):
            u, b = self._cache[len_cache - k]
            psif = linear_combination(1.0, psif, vf[N - k] / b, u)
        self._cache = []
        psif = self._rebuild(psif, N - len_cache - 1)
        nrm = norm(psif)
        self.host_syncs += 1
        return scalar_multiply(1.0 / nrm, psif)

    def run(self):
        N = self._build_krylov()
        E0 = float(self.Es[N - 1, 0])
        if self.E_shift is not None:
            E0 -= float(self.E_shift)
        if N == 1:
            return E0, self.psi0, N
        return E0, self._calc_result_full(N), N


class LanczosEvolution(LanczosGroundState):
    """``psi(t) = exp(delta H) psi0`` in the Krylov space (krylov_based.cpp:948-1024): the same recurrence and vector
    operations; only the small host algebra differs (exp of the tridiagonal matrix instead of its lowest eigenvector,
    convergence = weight of the last Krylov vector below ``P_tol``).  ``run(delta, normalize=None)`` returns
    ``(psi, N)``; ``normalize`` defaults to True for purely imaginary ``delta`` (real time).  For a real ``delta`` the
    coefficients are real and the result keeps the dtype of ``psi0``."""

    def __init__(self, H, psi0, options=None):
        super().__init__(H, psi0, options)
        self._result_norm = 1.0
        self.delta = None

    def run(self, delta, normalize=None):
        self.delta = complex(delta)
        self._cache = []
        self.h_krylov[:] = 0.0
        self.Es[:] = 0.0
        N = self._build_krylov()
        psi = scalar_multiply(self._coef(self._result_krylov[0]), self.psi0) if N == 1 else self._calc_result_full(N)
        if normalize if normalize is not None else (self.delta.real == 0.0):
            return psi, N
        return scalar_multiply(self._psi0_norm * self._result_norm, psi), N

    def _coef(self, c):
        return float(np.real(c)) if self.delta.imag == 0.0 else complex(c)

    def _calc_result_krylov(self, k):
        if k == 0:
            e = np.exp(self.delta * self.h_krylov[0, 0])
            self._result_norm = abs(e)
            self._result_krylov = np.asarray([e / self._result_norm])
        else:
            E, V = np.linalg.eigh(self.h_krylov[:k + 1, :k + 1])
            self.Es[k, :k + 1] = E
            x = V @ (np.exp(E * self.delta) * np.conj(V[0]))
            self._result_norm = float(np.linalg.norm(x))
            self._result_krylov = x / self._result_norm
        if self.delta.imag == 0.0:
            self._result_krylov = np.real(self._result_krylov)

    def _converged(self, k) -> bool:
        return abs(self._result_krylov[k]) < self.P_tol


def lanczos(H, psi, options=None):
    """``cyten.krylov_based.lanczos``: ``(E0, psi0, N)``."""
    return LanczosGroundState(H, psi, options).run()

#!/usr/bin/env python3
"""Benchmark of the abelian block-sparse hot path (BASELINE.json metric:
"U(1) tdot GFLOP/s and block truncated-SVD ms at chi=4096").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--chi 4096]

One *step* = one pass of the hot path over one synthetic input: the two-site DMRG effective-Hamiltonian
matvec  LP.W.W.RP.theta  (toycodes/tenpy_toycodes/d_dmrg.py:70-88: 5 permute_legs + 4 compose, i.e.
4 tdot-style contractions) on random U(1) tensors with the bond leg of BASELINE.md section 3.
``value`` = algorithmic flops of the 4 composes (sum 2*M*K*N over the reference's matrix_dot list,
abelian.cpp:1424-1446) / device time of the whole step, inputs resident in HBM.  The truncated SVD of
theta at the same chi is timed separately and reported as ``svd_ms``.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path


def _host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# The CPU legs (``--impl reference`` and the ``cpu_baseline`` of the b200 arm) use every host core the process may
# run on.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers, and OpenBLAS reads the variable when numpy
# is imported: so the override has to happen HERE, before the first ``import numpy``.
if "reference" in sys.argv[1:] or any(a.startswith("--impl=reference") for a in sys.argv[1:]):
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(_host_cores())

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "u1_tdot_gflops"
UNIT = "GFLOP/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--chi", type=int, default=4096)
    ap.add_argument("--svd-chi", type=int, default=None)
    ap.add_argument("--no-svd", action="store_true")
    ap.add_argument("--no-lanczos", action="store_true")
    ap.add_argument("--unfused", action="store_true",
                    help="time the call-for-call mirror of HEffective.matvec (5 permute_legs + 4 compose) instead of the fused "
                         "GEMM / one streaming pass / GEMM operator")
    ap.add_argument("--no-sweep", action="store_true", help="skip the sweep-step leg (BASELINE config 5)")
    ap.add_argument("--no-plugin", action="store_true", help="skip the e2e_plugin leg (cyten API through the plugin)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying a CUDA graph")
    ap.add_argument("--cpu-chi", type=int, default=None, help="bond dimension of the bounded CPU sample")
    return ap.parse_args()


# ---- clocks sampling ---------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def ncu_traffic(kernel):
    """DRAM bytes per launch of ``kernel`` from the newest committed ncu --set full capture
    (profiles/<round>/traffic.json), or None."""
    files = sorted((ROOT / "profiles").glob("*/traffic.json"))
    if not files:
        return None, None
    try:
        d = json.loads(files[-1].read_text()).get(kernel)
        return (d["bytes"], str(files[-1].relative_to(ROOT))) if d else (None, None)
    except Exception:
        return None, None


def measured_peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        d = json.loads(f.read_text())
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


# ---- reference arm: the UNMODIFIED reference (oracle/_ref) on the host cores ------------------------------
def build_reference_heff(chi, plugin=None):
    """H_eff and theta as REFERENCE tensors (cyten's own SymmetricTensor / permute_legs / compose) from the same host
    blocks as the native arm.  ``plugin``: a ``CytenB200Backend`` -- the tensors are then moved onto it, so the very same
    cyten calls run on the B200 kernels (the ``e2e_plugin`` leg)."""
    if str(ROOT / "oracle" / "_ref") not in sys.path:
        sys.path.insert(0, str(ROOT / "oracle" / "_ref"))
    import cyten as ct
    from cyten_b200.workloads import HEFF_LABELS, HEFF_SEEDS, HEffective, heff_spaces, random_blocks
    be = ct.get_backend("abelian", "numpy")
    sym = ct.u1_symmetry

    def conv(leg):
        return ct.ElementarySpace.from_defining_sectors(sym, leg.defining_sectors, leg.multiplicities, is_dual=leg.is_dual)
    T = {}
    for name, (cod, dom) in heff_spaces(chi).items():
        bi, blocks = random_blocks(cod, dom, HEFF_SEEDS[name])
        data = ct.backends.AbelianBackendData(dtype=ct.Dtype.float64, device="cpu",
                                              blocks=[be.block_backend.as_block(b) for b in blocks],
                                              block_inds=ct.backends.BlockInds(bi) if hasattr(ct.backends, "BlockInds") else bi,
                                              is_sorted=True)
        T[name] = ct.SymmetricTensor(data, [conv(l) for l in cod], [conv(l) for l in dom], backend=be,
                                     labels=HEFF_LABELS[name])
        if plugin is not None:
            if str(ROOT / "integration") not in sys.path:
                sys.path.insert(0, str(ROOT / "integration"))
            from cyten_plugin import to_plugin
            T[name] = to_plugin(T[name], plugin)

    class Ops:
        permute_legs = staticmethod(lambda t, codomain=None, domain=None: ct.permute_legs(t, codomain, domain, bend_right=None))
        compose = staticmethod(ct.compose)
    return HEffective(T["LP"], T["RP"], T["W1"], T["W2"], Ops), T["theta"], ct


class ReferenceOps:
    """The ``SweepStep`` operation interface (cyten_b200.workloads.NativeOps) over the reference's own functions."""

    def __init__(self, ct):
        self.ct = ct

    @staticmethod
    def relabel(T, mapping):
        T = T.copy(deep=False)
        T.relabel({k: v for k, v in mapping.items() if k in T.labels})
        return T

    def permute_legs(self, T, codomain=None, domain=None):
        return self.ct.permute_legs(T, codomain, domain, bend_right=None)

    def compose(self, A, B, relabel1=None, relabel2=None):
        return self.ct.compose(A, B, relabel1=relabel1, relabel2=relabel2)

    def truncated_svd(self, theta, new_labels, chi_max, svd_min):
        U, S, Vh, err, _ = self.ct.truncated_svd(theta, new_labels, chi_max=chi_max, svd_min=svd_min)
        return U, S, Vh, err

    def scale_axis(self, T, D, leg):
        return self.ct.scale_axis(T, D, leg=leg)

    def pinv(self, D, cutoff):
        return self.ct.pinv(D, cutoff=cutoff)

    def hc(self, T):
        return T.hc

    def norm(self, T):
        x = self.ct.norm(T)
        return float(x.to_numpy()) if hasattr(x, "to_numpy") else float(x)

    def scale(self, T, a):
        return a * T

    @staticmethod
    def shard(T, leg):
        return T

    @staticmethod
    def gather(T):
        return T


def build_reference_sweep_step(chi, plugin=None, nmv=1, chi_max=None, eps=1e-14, sync=None):
    """``SweepStep`` (BASELINE config 5) on reference tensors -- numpy backend, or the plugin if given -- from the same
    host blocks and Schmidt values as cyten_b200.workloads.build_sweep_step."""
    from cyten_b200.workloads import SweepStep, heff_spaces, schmidt_values
    H, theta, ct = build_reference_heff(chi, plugin=plugin)
    o = ReferenceOps(ct)
    LP = o.permute_legs(H.LP, ["vL"], ["vL*", "wL*"])
    RP = o.permute_legs(H.RP, ["vR"], ["wR*", "vR*"])
    W = o.permute_legs(o.relabel(H.W1, {"p0": "p", "p0*": "p*", "wC": "wR"}), ["wL", "p"], ["p*", "wR"])
    leg = theta.codomain.factors[0]
    vals = schmidt_values(heff_spaces(chi)["theta"][0][0])
    be = theta.backend
    S = ct.DiagonalTensor.from_diag_block(be.block_backend.as_block(np.ascontiguousarray(vals)), leg, backend=be, labels=["vL", "vR"])
    if plugin is not None:
        from cyten_plugin import to_plugin
        S = to_plugin(ct.DiagonalTensor.from_diag_block(ct.get_backend("abelian", "numpy").block_backend.as_block(
            np.ascontiguousarray(vals)), leg, backend=ct.get_backend("abelian", "numpy"), labels=["vL", "vR"]), plugin)
    return SweepStep(o, LP, RP, W, S, chi_max or chi, eps, nmv, sync), theta, ct


def cpu_matvec_time(chi, kind, repeats=1):
    """Seconds per matvec of the CPU implementation (kind 'reference' = oracle/_ref, 'port' = numpy oracle)."""
    if kind == "reference":
        H, theta, _ = build_reference_heff(chi)
    else:
        from cyten_b200.backend import B200AbelianBackend
        from cyten_b200.workloads import build_heff
        from oracle import cabi_numpy
        H, theta, _, _ = build_heff(chi, B200AbelianBackend("cpu", _library=cabi_numpy))
    H.matvec(theta)   # warm-up
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        H.matvec(theta)
        best = min(best, time.perf_counter() - t0)
    return best


def build_reference_theta(chi):
    """theta[vL, p0 | p1, vR] with N(0,1) blocks (from_random_normal on the two-site legs) on the
    reference's numpy backend -- the same host blocks as the GPU arm (workloads.random_blocks, seed 9)."""
    sys.path.insert(0, str(ROOT / "oracle" / "_ref"))
    import cyten as ct
    from cyten_b200.workloads import CONFIG_LEGS, random_blocks, u1_bond_leg, u1_phys_leg
    be = ct.get_backend("abelian", "numpy")
    sym = ct.u1_symmetry
    Q, sigma = CONFIG_LEGS[chi]
    v, p = u1_bond_leg(chi, Q, sigma), u1_phys_leg()

    def conv(leg):
        return ct.ElementarySpace.from_defining_sectors(sym, leg.defining_sectors, leg.multiplicities, is_dual=leg.is_dual)
    bi, blocks = random_blocks([v, p], [p.dual, v], 9)
    data = ct.backends.AbelianBackendData(dtype=ct.Dtype.float64, device="cpu",
                                          blocks=[be.block_backend.as_block(b) for b in blocks],
                                          block_inds=ct.backends.BlockInds(bi), is_sorted=True)
    theta = ct.SymmetricTensor(data, [conv(v), conv(p)], [conv(p.dual), conv(v)], backend=be,
                               labels=["vL", "p0", "p1", "vR"])
    return theta, ct


def cpu_svd_time(chi, kind):
    """Seconds for one truncated_svd(theta, chi_max=chi, svd_min=1e-14) on the host cores."""
    if kind == "reference":
        theta, ct = build_reference_theta(chi)
        t0 = time.perf_counter()
        ct.truncated_svd(theta, chi_max=chi, svd_min=1e-14)
        return time.perf_counter() - t0
    from cyten_b200 import tensors as ops
    from cyten_b200.backend import B200AbelianBackend
    from cyten_b200.workloads import CONFIG_LEGS, mps_tensor, u1_bond_leg, u1_phys_leg
    from oracle import cabi_numpy
    be = B200AbelianBackend("cpu", _library=cabi_numpy)
    Q, sigma = CONFIG_LEGS[chi]
    v, p = u1_bond_leg(chi, Q, sigma), u1_phys_leg()
    from cyten_b200.workloads import random_tensor
    th = random_tensor([v, p], [p.dual, v], be, seed=9, labels=["vL", "p0", "p1", "vR"])
    t0 = time.perf_counter()
    ops.truncated_svd(th, chi_max=chi, svd_min=1e-14)
    return time.perf_counter() - t0


def blas_threads(want: int) -> int:
    """Sets the BLAS/OpenMP pools of this process to ``want`` threads (works after numpy was imported, unlike the
    environment variables) and returns the number of threads the BLAS behind numpy really uses."""
    try:
        import threadpoolctl
        threadpoolctl.threadpool_limits(limits=want)
        pools = [p.get("num_threads", 1) for p in threadpoolctl.threadpool_info() if p.get("user_api") == "blas"]
        return int(max(pools)) if pools else want
    except Exception:   # noqa: BLE001
        return want


def reference_available():
    return bool(list((ROOT / "oracle" / "_ref" / "cyten").glob("_core*.so")))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind = "reference" if reference_available() else "port"
    chi = args.chi
    cores = blas_threads(_host_cores())
    flops = flops_of_matvec(chi)
    if kind == "reference":
        H, theta, _ = build_reference_heff(chi)
    else:
        from cyten_b200.backend import B200AbelianBackend
        from cyten_b200.workloads import build_heff
        from oracle import cabi_numpy
        H, theta, _, _ = build_heff(chi, B200AbelianBackend("cpu", _library=cabi_numpy))
    warm = max(1, min(args.warmup, 5))
    for _ in range(warm):
        H.matvec(theta)
    steps = max(1, min(args.steps, 100))     # 0.3 s per matvec at chi=4096 on 16 cores: K steps stay within minutes
    t0 = time.perf_counter()
    for _ in range(steps):
        H.matvec(theta)
    dt = (time.perf_counter() - t0) / steps
    val = flops / dt / 1e9
    svd_ms = None
    if not args.no_svd:
        svd_ms = cpu_svd_time(args.svd_chi or chi, kind) * 1e3
    lanczos_info = None
    if kind == "reference" and not args.no_lanczos and not args.no_svd:
        # the loop the matvec sits in: the reference's own Lanczos, same operator, same 10 forced iterations as the
        # b200 arm's `lanczos` key
        try:
            from cyten import krylov_based, sparse

            class RefOp(sparse.LinearOperator):
                def matvec(self, t):
                    return H.matvec(t)
            t0 = time.perf_counter()
            _, _, nit = krylov_based.lanczos(RefOp(), theta, {"N_max": 10, "N_min": 10, "P_tol": 0.0})
            ms = (time.perf_counter() - t0) * 1e3
            lanczos_info = {"iterations": int(nit), "ms": ms, "ms_per_iteration": ms / int(nit),
                            "what": "cyten.krylov_based.lanczos on the host cores"}
        except Exception as e:   # noqa: BLE001
            lanczos_info = {"error": repr(e)}
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "impl": "reference", "lanczos": lanczos_info,
            "config": {"workload": f"U(1) DMRG two-site H_eff matvec (4 tdot + 5 permute_legs), chi={chi}, D=5, d=2, float64"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{steps} matvec(s) at chi={chi} on the host cores"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "svd_ms": svd_ms}
    print(json.dumps(line))


_FLOPS_CACHE = {}


def flops_of_matvec(chi):
    """2 * sum M*K*N over the matrix_dot lists of the 4 composes (abelian.cpp:1424-1446).  Depends on
    block_inds + multiplicities only: a plan-only (dry_run) pass of the host logic over the numpy oracle."""
    if chi in _FLOPS_CACHE:
        return _FLOPS_CACHE[chi]
    import torch
    from cyten_b200 import tensors as ops
    from cyten_b200.backend import B200AbelianBackend, conventional_leg_order
    from cyten_b200.data import AbelianBackendData
    from cyten_b200.spaces import TensorProduct
    from cyten_b200.tensors import SymmetricTensor, valid_block_inds
    from cyten_b200.workloads import HEFF_LABELS, HEffective, heff_spaces
    from oracle import cabi_numpy
    be = B200AbelianBackend("cpu", _library=cabi_numpy)
    be.dry_run = True
    be.plan_mkn = []
    T = {}
    for name, (cod, dom) in heff_spaces(chi).items():
        cod, dom = TensorProduct(cod), TensorProduct(dom)
        bi = valid_block_inds(cod, dom)
        legs = conventional_leg_order(cod, dom)
        shapes = np.stack([np.asarray(l.multiplicities)[bi[:, ax]] for ax, l in enumerate(legs)], axis=1)
        T[name] = SymmetricTensor(be._empty("float64", "cpu", bi, shapes), cod, dom, be, HEFF_LABELS[name])
    HEffective(T["LP"], T["RP"], T["W1"], T["W2"], ops).matvec(T["theta"])
    _FLOPS_CACHE[chi] = 2.0 * float(sum(be.plan_mkn))
    return _FLOPS_CACHE[chi]


# ---- the B200 arm -------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from cyten_b200 import tensors as ct
    from cyten_b200 import sharding as sh
    from cyten_b200.backend import B200AbelianBackend
    from cyten_b200.workloads import build_heff, CONFIG_LEGS

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    be = B200AbelianBackend(dev)
    pg = sh.attach_process_group(be, dist.group.WORLD if world > 1 else None)
    chi = args.chi
    H_unfused, theta_full, host, sp = build_heff(chi, be)
    # the product path: GEMM, one fused streaming pass for the two MPO applications, GEMM (cyten_b200/fused.py); --unfused
    # times the call-for-call mirror of the reference's 5 permute_legs + 4 compose instead.  `flops` is always the
    # algorithmic GEMM work of the reference's four composes (same number for both, and for the reference arm).
    fused = not args.unfused
    if fused:
        from cyten_b200.fused import FusedHEffective
        H = FusedHEffective(H_unfused)
    else:
        H = H_unfused
    SH = sh.ShardedHEffective(H_unfused, pg, split_columns=os.environ.get("CYB_SPLIT_COLUMNS", "1") == "1",
                              column_heavy=float(os.environ.get("CYB_COL_HEAVY", "0.35")), fused=fused)
    # multi-GPU: theta lives sharded along the sectors of vR (LPT by block volume); the matvec chain is
    # rank-local until vR is contracted, then one all_reduce of the (small) result arena
    theta = SH.shard(theta_full) if world > 1 else theta_full

    # N > 1, fused operator, CYB_OVERLAP=1: the last GEMM runs in column groups of the result and each group's all_reduce is
    # issued as soon as the group is queued (cyten_b200.sharding.OverlappedMatvec).  Measured SLOWER on 2 B200 (1.24 vs
    # 1.11 ms, profiles/r2/scale/r2t_*): the persistent GEMM CTAs fill every SM (4 x 56 KB of shared memory), so NCCL's
    # CTAs only become resident when the GEMM drains and the second launch costs more than the collective it was to hide.
    # Default: whole chain, then one all_reduce.
    overlap = world > 1 and fused and os.environ.get("CYB_OVERLAP", "0") == "1"
    overlap_fractions = tuple(float(x) for x in os.environ.get("CYB_OVERLAP_FRACTIONS", "0.6,0.4").split(","))
    overlapped = {}

    def ov_of(t):
        if id(t) not in overlapped:
            overlapped[id(t)] = sh.OverlappedMatvec(SH, t, fractions=overlap_fractions)
        return overlapped[id(t)]

    def eager(t):
        if overlap:
            return ov_of(t)(t)
        return SH.matvec(t) if world > 1 else H.matvec(t)

    def local_chain(t):      # the part of a step without a collective: what a CUDA graph captures
        # (the all_reduce as a node of the same graph was measured on 2 B200: 1.105 ms against 1.103 ms, and the process group
        #  did not shut down cleanly afterwards -- the collective stays an eager launch between the replays)
        return SH.matvec_local(t) if world > 1 else H.matvec(t)
    captured = {}

    def capture(t):
        if overlap:
            from cyten_b200.graphs import CapturedPipeline
            ov = ov_of(t)
            return CapturedPipeline([lambda x, g=g: ov.part(g, x) for g in range(ov.num_parts)], t, warmup=2)
        from cyten_b200.graphs import CapturedStep
        return CapturedStep(local_chain, t, warmup=2)

    def step(t=None):
        t = theta if t is None else t
        cap = captured.get(id(t))
        if cap is None:
            return eager(t)
        if overlap:
            return ov_of(t)(t, run_part=cap.replay)
        y = cap()
        if world > 1:
            pg.all_reduce(y.data.buf)
        return y

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def total_over_ranks(x, op=None):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=op or dist.ReduceOp.SUM)
        return float(t.item())

    # cold pass (plans are built on device here) -- timed separately, reported as ms_first_call
    barrier()
    t0 = time.perf_counter()
    y = step()
    torch.cuda.synchronize()
    first_call_ms = (time.perf_counter() - t0) * 1e3
    # algorithmic flops of the whole job = GEMM work of the reference's call sequence (4 composes) on this structure,
    # counted by the planner during one unsharded, unfused matvec on this GPU
    be.plan_mkn = []
    H_unfused.matvec(theta_full).data.buf
    flops = 2.0 * float(sum(be.plan_mkn))
    be.plan_mkn = None
    for _ in range(max(args.warmup, 3)):
        y = step()
    be.plan_mkn = None
    l2_flush = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)
    # the plan-cached chain is pure kernel launches: capture it once into a CUDA graph and replay it
    use_graph = not args.no_graph
    if use_graph:
        captured[id(theta)] = capture(theta)
        for _ in range(2):
            y = step()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # ---- timed region: K steps, device time, L2 flushed between steps (flush excluded via events) ----
    def timed_steps(fn, nsteps):
        evs = []
        barrier()
        for _ in range(nsteps):
            l2_flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        barrier()
        return sum(a.elapsed_time(b) for a, b in evs)
    be.launches = 0
    total_ms = timed_steps(step, args.steps)
    launches = be.launches
    # per-launch timing of the GEMM kernels (roofline): the same kernels launched eagerly with CUDA events around
    # each launch on the launching stream (events cannot be recorded inside a graph replay)
    be.gemm_events = []
    be.gemm_event_flops = []
    eager_total_ms = timed_steps(lambda: eager(theta), args.steps)
    gemm_ms = [a.elapsed_time(b) for a, b in be.gemm_events]
    gemm_flops = list(be.gemm_event_flops)
    be.gemm_events = None
    eager_ms_per_step = total_over_ranks(eager_total_ms, dist.ReduceOp.MAX if world > 1 else None) / args.steps
    total_ms = total_over_ranks(total_ms, dist.ReduceOp.MAX if world > 1 else None)
    ms_per_step = total_ms / args.steps
    value = flops / (ms_per_step * 1e-3) / 1e9

    # ---- N > 1: the WHOLE reduced result against the unsharded operator on this GPU (outside every timed region) ----
    sharded_check = None
    if world > 1:
        y_sh = step()
        y_un = sh.expand_to_full(H.matvec(theta_full))
        same_blocks = bool(np.array_equal(y_sh.data.block_inds, y_un.data.block_inds))
        # block by block: the padding between blocks is not part of the tensor (and not zeroed when a result is complete already)
        nblk = y_un.data.num_blocks if same_blocks else 0
        scale = max((float(y_un.data.block(i).abs().max()) for i in range(nblk)), default=0.0)
        diff = (max((float((y_sh.data.block(i) - y_un.data.block(i)).abs().max()) for i in range(nblk)), default=0.0)
                if same_blocks else float("inf"))
        worst = total_over_ranks(diff / scale if scale > 0 else diff, dist.ReduceOp.MAX)
        sharded_check = {"max_abs_diff_over_max_abs": worst, "tolerance": 1e-12, "ok": bool(same_blocks and worst <= 1e-12),
                         "what": "graph-replayed sharded step (all ranks reduced) vs the unsharded fused operator, every element"}

    # ---- the same step with the plan cache disabled (bookkeeping recomputed on device every call) ----
    be.plan_cache = False
    eager(theta)
    nnp = min(args.steps, 3)
    noplan_ms = total_over_ranks(timed_steps(lambda: eager(theta), nnp) / nnp, dist.ReduceOp.MAX if world > 1 else None)
    be.plan_cache = True

    # ---- end to end: theta from pinned host memory every step, result back to the host ----
    # (1) serial: upload -> matvec -> download -> synchronize, one step after the other (latency);
    # (2) pipelined (the e2e value): the same three stages on three streams with two theta arenas and
    #     two host result buffers, so the upload of step i+1 and the download of step i-1 overlap the
    #     kernels of step i.  Every step still moves its own input and its own result inside the timed
    #     region.  With N ranks every rank uploads its shard of theta and downloads 1/N of the result.
    from cyten_b200.data import AbelianBackendData
    from cyten_b200.tensors import SymmetricTensor
    nel = theta.data.buf.numel()
    pinned = torch.empty(nel, dtype=torch.float64).pin_memory()
    pinned.copy_(theta.data.buf)
    d2 = AbelianBackendData(theta.data.dtype, dev, torch.empty_like(theta.data.buf), theta.data.block_inds,
                            theta.data.offsets, theta.data.shapes, sig=theta.data.sig)
    thetas = [theta, SymmetricTensor(d2, theta.codomain, theta.domain, be, theta.labels)]

    if use_graph:
        captured[id(thetas[1])] = capture(thetas[1])   # one graph (or pipeline of graphs) per theta arena
    stepk = step
    y = stepk(thetas[1])
    nres = y.data.buf.numel()
    lo, hi = (rank * nres) // world, ((rank + 1) * nres) // world
    out_host = [torch.empty(hi - lo, dtype=torch.float64).pin_memory() for _ in range(2)]
    h2d, d2h = nel * 8, (hi - lo) * 8
    e2e_steps = max(1, min(args.steps, 10))
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        theta.data.buf.copy_(pinned, non_blocking=True)
        y = stepk(theta)
        out_host[0].copy_(y.data.buf[lo:hi], non_blocking=True)
        torch.cuda.synchronize()
    barrier()
    e2e_serial_s = total_over_ranks((time.perf_counter() - t0) / e2e_steps, dist.ReduceOp.MAX if world > 1 else None)

    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    cur = torch.cuda.current_stream(dev)
    pipe_steps = max(4, 2 * e2e_steps)

    def pipelined(nsteps):
        ev_comp, ev_out, keep = [None, None], [None, None], [None, None]
        for i in range(nsteps):
            k = i & 1
            if ev_comp[k] is not None:
                s_in.wait_event(ev_comp[k])          # step i-2 has consumed arena k
            with torch.cuda.stream(s_in):
                thetas[k].data.buf.copy_(pinned, non_blocking=True)
                e_in = torch.cuda.Event()
                e_in.record(s_in)
            cur.wait_event(e_in)
            if ev_out[k] is not None:
                cur.wait_event(ev_out[k])            # a graph's result arena is static: step i-2's download must be done
            yk = stepk(thetas[k])
            yb = yk.data.buf                         # materialises the final (lazy) permute_legs
            ec = torch.cuda.Event()
            ec.record(cur)
            ev_comp[k] = ec
            if ev_out[k] is not None:
                ev_out[k].synchronize()              # host buffer k and the result arena of step i-2 are free
            s_out.wait_event(ec)
            with torch.cuda.stream(s_out):
                out_host[k].copy_(yb[lo:hi], non_blocking=True)
                eo = torch.cuda.Event()
                eo.record(s_out)
            ev_out[k] = eo
            keep[k] = yk
        torch.cuda.synchronize()
    pipelined(2)
    barrier()
    t0 = time.perf_counter()
    pipelined(pipe_steps)
    barrier()
    e2e_s = total_over_ranks((time.perf_counter() - t0) / pipe_steps, dist.ReduceOp.MAX if world > 1 else None)
    e2e_val = flops / e2e_s / 1e9
    e2e_check = float(out_host[(pipe_steps - 1) & 1][:8].abs().sum())   # the result really arrived on the host

    # ---- the same step through the reference-facing plugin: cyten.permute_legs / cyten.compose on tensors that live on
    # CytenB200Backend (integration/cyten_plugin.py), theta uploaded from pinned host memory and the result downloaded
    # every step, serial (no pipelining, no CUDA graph: what a cyten user gets by switching the backend) ----
    plugin_info = None
    if world == 1 and not args.no_plugin and reference_available():
        try:
            plugin_info = e2e_plugin(args, chi, dev, flops, y)
        except Exception as e:   # noqa: BLE001 -- an extra, must never cost the bench line
            plugin_info = {"error": repr(e)}

    # ---- BASELINE config 5: one bond update of a DMRG sweep (matvec + truncated_svd + scale_axis/pinv + update_LP/RP) ----
    sweep_info = None
    if world == 1 and not args.no_sweep and not args.no_svd:
        try:
            sweep_info = sweep_bench(chi, be)
        except Exception as e:   # noqa: BLE001 -- an extra, must never cost the bench line
            sweep_info = {"error": repr(e)}

    # ---- truncated SVD of the two-site theta at chi (BASELINE config 3) ----
    # theta[vL, p0 | p1, vR] with N(0,1) blocks (the synthetic input of SURVEY 8(d)); the rank-chi theta = tdot(A, B)
    # of two MPS tensors (half of every sector's singular values are numerically zero, the slow case for a Jacobi
    # method) is timed too and reported as svd.rank_deficient_ms.
    svd_ms = svd_info = None
    if not args.no_svd:
        from cyten_b200.workloads import mps_tensor, random_tensor, u1_bond_leg, u1_phys_leg
        schi = args.svd_chi or chi
        Q, sigma = CONFIG_LEGS[schi]
        v, p = u1_bond_leg(schi, Q, sigma), u1_phys_leg()
        th_full = random_tensor([v, p], [p.dual, v], be, seed=9, labels=["vL", "p0", "p1", "vR"])
        th_def = ct.tdot(mps_tensor(v, p, v, be, seed=7), mps_tensor(v, p, v, be, seed=8), [2], [0])
        th_def.data.buf

        def time_svd(th, reps):
            def do_svd():
                if world > 1:
                    return sh.truncated_svd_sharded(th, pg, chi_max=schi, svd_min=1e-14)[:5]
                return ct.truncated_svd(th, chi_max=schi, svd_min=1e-14)
            do_svd()   # warm-up
            times = []
            for _ in range(reps):
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                U, S, Vh, err, _ = do_svd()
                e1.record()
                barrier()
                times.append(e0.elapsed_time(e1))
            ms = total_over_ranks(min(times), dist.ReduceOp.MAX if world > 1 else None)
            kept = int(total_over_ranks(float(S.data.shapes.sum()))) if world > 1 else int(S.data.shapes.sum())
            sweeps = int(be.last_info.abs().max().item()) if be.last_info is not None and be.last_info.numel() else None
            return ms, kept, err, sweeps
        svd_ms, kept, err, sweeps = time_svd(th_full, 3)
        def_ms, def_kept, def_err, def_sweeps = time_svd(th_def, 2)
        svd_info = {"chi": schi, "input": "N(0,1) blocks on [vL,p0 | p1,vR] (full rank)", "kept": kept, "err": err,
                    "sweeps_max": sweeps, "rank_deficient_ms": def_ms, "rank_deficient_kept": def_kept,
                    "rank_deficient_sweeps_max": def_sweeps}
    # ---- the loop the matvec sits in: 10 Lanczos iterations on this H_eff, graph-replayed matvec,
    # device-resident alpha, one read-back per iteration.  The random environments are not Hermitian, so the Ritz value
    # means nothing: a timing; parity of the loop is tests/test_krylov.py and tests/test_live_reference.py.
    lanczos_info = None
    if world == 1 and not args.no_lanczos and not args.no_svd:
        try:
            from cyten_b200.krylov import GraphedOperator, LanczosGroundState
            G = GraphedOperator(H)
            iters = 10
            lopts = {"N_max": iters, "N_min": iters, "P_tol": 0.0}
            LanczosGroundState(G, theta_full, dict(lopts)).run()          # warm-up: plans + graph capture
            best = float("inf")
            for _ in range(3):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                lz = LanczosGroundState(G, theta_full, dict(lopts))
                E0, psi, nit = lz.run()
                psi.data.buf
                torch.cuda.synchronize()
                best = min(best, (time.perf_counter() - t0) * 1e3)
            lanczos_info = {"iterations": int(nit), "ms": best, "ms_per_iteration": best / nit, "host_syncs": lz.host_syncs,
                            "what": "LanczosGroundState(GraphedOperator(H_eff), theta): wall time incl. result vector"}
        except Exception as e:   # noqa: BLE001 -- an extra, must never cost the bench line
            lanczos_info = {"error": repr(e)}
    clocks = sampler.stop() if rank == 0 else None

    if rank == 0:
        # ---- roofline of the dominant kernel (grouped DMMA GEMM): FP64 tensor peak ----
        # per-launch mean over the launches that carry > 5 % of the step's flops (the tensor-core
        # GEMMs; the MPO applications are HBM-bound and run on the streaming kernel)
        fat = [(m, f) for m, f in zip(gemm_ms, gemm_flops) if f > 0.05 * max(gemm_flops)] if gemm_ms else []
        peak_tf, peak_src = fp64_peak_tflops(dev)
        roof = None
        if fat:
            g_ms = float(np.mean([m for m, _ in fat]))
            g_fl = float(np.mean([f for _, f in fat]))
            ach = g_fl / (g_ms * 1e-3) / 1e12
            traffic, traffic_src = ncu_traffic("grouped_gemm_kernel") if (world == 1 and chi == 4096) else (None, None)
            roof = {"bound": "tensor", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf,
                    "traffic": traffic, "traffic_source": traffic_src,
                    "kernel": "grouped_gemm_kernel (DMMA f64)", "peak_source": peak_src,
                    "launch_ms": g_ms, "launch_flops": g_fl,
                    "gemm_share_of_step": sum(gemm_ms) / eager_total_ms if eager_total_ms else None}
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            cpu = cpu_baseline(args)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": f"U(1) DMRG two-site H_eff matvec (4 tdot + 5 permute_legs), chi={chi}, D=5, d=2, float64",
                           "implementation": ("fused: GEMM, one cyb_middle_apply pass for both MPO applications, GEMM (cyten_b200/fused.py)"
                                              if fused else "unfused: 5 permute_legs + 4 compose, call for call"),
                           "l2": "192 MiB buffer written between timed steps", "flops_per_step": flops,
                           "cuda_graph": {"enabled": use_graph, "ms_per_step_eager": eager_ms_per_step,
                                          "launches_per_replay": (captured[id(theta)].launches_per_replay if use_graph else None)},
                           "plan_cache": {"enabled": True, "hits": be.plan_hits, "misses": be.plan_misses,
                                          "ms_per_step_without": noplan_ms, "ms_first_call": first_call_ms},
                           "sharding": (("theta sharded along vR (sectors and index ranges, LPT by flops), rank-local chain; last GEMM in "
                                         f"{len(overlap_fractions)} column groups {overlap_fractions}, each group's all_reduce behind the next group's GEMM")
                                        if overlap else
                                        "theta sharded along vR sectors (LPT), rank-local chain, one all_reduce per matvec"
                                        if world > 1 else "single GPU")},
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": e2e_s * 1e3, "ms_per_step_serial": e2e_serial_s * 1e3, "steps": pipe_steps,
                        "mode": "3-stream pipeline (upload | matvec | download), two theta arenas; bytes are per rank",
                        "host_checksum": e2e_check},
                "gpu_launches": launches, "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
                "e2e_plugin": plugin_info, "sweep": sweep_info, "svd_ms": svd_ms, "svd": svd_info, "lanczos": lanczos_info,
                "sharded_check": sharded_check}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def sweep_bench(chi, be, reps=3):
    """One ``update_bond`` of the reference's toy DMRG (cyten_b200.workloads.SweepStep; d_dmrg.py:217-274) on the chi legs:
    device time of the whole step (CUDA events, best of ``reps``) and the share of each part (wall clock with a
    synchronisation after every part, second pass)."""
    import torch
    from cyten_b200.workloads import build_sweep_step
    st, th = build_sweep_step(chi, be, fused=True)
    st.update_bond(th)                       # plans
    best = float("inf")
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = st.update_bond(th)
        r["LP"].data.buf, r["RP"].data.buf   # materialise the final (deferred) permutations
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    st.sync = torch.cuda.synchronize
    st.timings = {}
    st.update_bond(th)
    tot = sum(st.timings.values())
    return {"chi": chi, "ms": best, "parts_ms": {k: v * 1e3 for k, v in st.timings.items()},
            "share": {k: v / tot for k, v in st.timings.items()}, "kept": int(r["S"].data.shapes.sum()), "err": r["err"],
            "what": "update_bond: 4 bends + 1 H_eff matvec + normalise, truncated_svd(chi_max=chi, svd_min=1e-14), "
                    "scale_axis/pinv, update_LP, update_RP"}


def e2e_plugin(args, chi, dev, flops, y_native):
    """One H_eff matvec per step written with cyten's OWN calls (toycodes/tenpy_toycodes/d_dmrg.py:70-88) on tensors of
    the plugin backend; every step uploads theta from pinned host memory, runs the 5 permute_legs + 4 compose virtuals
    (each one: reference C++ -> trampoline -> plugin -> cyb:: kernels) and downloads the result."""
    import torch
    if str(ROOT / "oracle" / "_ref") not in sys.path:
        sys.path.insert(0, str(ROOT / "oracle" / "_ref"))
    if str(ROOT / "integration") not in sys.path:
        sys.path.insert(0, str(ROOT / "integration"))
    from cyten_plugin import install
    plugin = install(str(dev), key="b200-bench")
    H, theta, ct = build_reference_heff(chi, plugin=plugin)
    arena = plugin._arena(theta.data, theta.num_legs)          # theta's device arena (its blocks are views into it)
    nel = arena.buf.numel()
    pinned = torch.empty(nel, dtype=torch.float64).pin_memory()
    pinned.copy_(arena.buf)
    y = H.matvec(theta)
    ya = plugin._arena(y.data, y.num_legs)
    out = torch.empty(ya.buf.numel(), dtype=torch.float64).pin_memory()
    # same host blocks, same kernels, same plans as the native arm: the results must agree
    yn = y_native.data
    same = bool(ya.num_blocks == yn.num_blocks and np.array_equal(ya.block_inds, yn.block_inds) and
                all(torch.allclose(ya.block(i), yn.block(i), rtol=0, atol=1e-10 * float(yn.block(i).abs().max()))
                    for i in range(yn.num_blocks)))
    for _ in range(2):
        H.matvec(theta)
    for c in (plugin.hot_calls, plugin.cold_calls, plugin.fallbacks):
        c.clear()
    plugin.fast.launches = 0
    steps = max(3, min(args.steps, 10))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        arena.buf.copy_(pinned, non_blocking=True)
        y = H.matvec(theta)
        out.copy_(plugin._arena(y.data, y.num_legs).buf, non_blocking=True)
        torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    return {"value": flops / dt / 1e9, "unit": UNIT, "ms_per_step": dt * 1e3, "steps": steps,
            "h2d_bytes_per_step": nel * 8, "d2h_bytes_per_step": out.numel() * 8,
            "api": "cyten.permute_legs / cyten.compose on CytenB200Backend tensors (get_backend('abelian', 'b200-bench'))",
            "hot_calls_per_step": {k: v / steps for k, v in plugin.hot_calls.items()},
            "fallbacks": dict(plugin.fallbacks), "cold_calls_per_step": {k: v / steps for k, v in plugin.cold_calls.items()},
            "kernel_launching_calls_per_step": plugin.fast.launches / steps, "matches_native_arm": same,
            "host_checksum": float(out[:8].abs().sum())}


def fp64_peak_tflops(dev):
    """FP64 tensor (DMMA) peak: MEASURED_PEAKS.json carries only bf16 + HBM, so the denominator is
    measured here with a cuBLAS DGEMM 8192^3 (best of 5, CUDA events) -- 'of measured (cuBLAS DGEMM)'."""
    import torch
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.matmul(a, b)
    best = float("inf")
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12, "cuBLAS DGEMM 8192^3 measured in this run"


def cpu_baseline(args):
    cores = blas_threads(_host_cores())
    kind = "reference" if reference_available() else "port"
    chi = args.cpu_chi or min(args.chi, 4096)
    try:
        dt = cpu_matvec_time(chi, kind)
    except Exception as e:   # the reference .so may not load on an odd box: fall back to the port
        kind = "port"
        dt = cpu_matvec_time(chi, kind)
    fl = flops_of_matvec(chi)
    svd_ms = None
    if not args.no_svd:
        try:
            svd_ms = cpu_svd_time(args.svd_chi or chi, kind) * 1e3
        except Exception as e:
            svd_ms = f"failed: {e!r}"
    return {"value": fl / dt / 1e9, "unit": UNIT, "cores": cores, "kind": kind, "ms_per_step": dt * 1e3, "svd_ms": svd_ms,
            "sample": f"1 H_eff matvec at chi={chi} (after 1 warm-up) + 1 truncated_svd(theta) at chi={args.svd_chi or chi}, "
                      "AbelianBackend+NumpyBlockBackend on the host cores"}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()

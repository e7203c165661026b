#!/usr/bin/env python3
"""Per-row measurement of the SURVEY section 8(a) hot-path operations on one B200.

  python tools/bench_ops.py [--chi 4096] [--dtype float64] [--no-ref] [--out gpurun_out/ops.json]

For every row (permute_legs, combine_legs, split_legs, tdot, svd, truncated_svd, qr, eigh) it reports
device time (CUDA events, best of ``--reps`` after warm-up, plan cache warm), the ALGORITHMIC work of
SURVEY section 8(d) (bytes for the copies, LAPACK-formula flops for the factorisations, sum 2MKN for
tdot), the resulting GB/s or TFLOP/s, and the time of the UNMODIFIED reference
(AbelianBackend+NumpyBlockBackend from oracle/_ref) on the host cores for the same inputs.
One JSON line per row, and a markdown table on stderr.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chi", type=int, default=4096)
    ap.add_argument("--dtype", default="float64", choices=["float64", "complex128"])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--ref-torch", action="store_true",
                    help="also time the reference's own GPU path (AbelianBackend + TorchBlockBackend on cuda: torch.matmul per "
                         "block pair, cuSOLVER per sector; src/block_backend/torch.cpp:1341-1394, 898-911) on the same inputs")
    ap.add_argument("--rows", default="permute,combine,split,tdot,svd,tsvd,qr,eigh,blas1")
    ap.add_argument("--out", default=None)
    ap.add_argument("--hubbard", action="store_true",
                    help="BASELINE config 4 legs: Fermi-Hubbard U(1) x U(1) (charge, 2Sz), d = 4, 2-D Gaussian multiplicities "
                         "(cyten_b200.workloads.hubbard_bond_leg): many small irregular blocks")
    return ap.parse_args()


def legs_for(chi, hubbard=False):
    from cyten_b200.workloads import CONFIG_LEGS, u1_bond_leg, u1_phys_leg
    if hubbard:
        from cyten_b200.workloads import hubbard_bond_leg, hubbard_phys_leg
        return hubbard_bond_leg(chi), hubbard_phys_leg()
    if chi in CONFIG_LEGS:
        Q, sigma = CONFIG_LEGS[chi]
    else:   # interpolate the Gaussian width with log2(chi) like the named configs
        Q = max(2, int(round(3 + 1.5 * (np.log2(chi) - 7))))
        sigma = 1.2 + 0.5 * (np.log2(chi) - 7)
    return u1_bond_leg(chi, Q, sigma), u1_phys_leg()


def gpu_time(fn, reps, sync_result):
    """Best-of-reps device time in ms; ``sync_result(res)`` forces deferred work (lazy permutes)."""
    import torch
    for _ in range(2):
        sync_result(fn())
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sync_result(fn())
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def touch(res):
    """Materialises every arena of a result (tuple of tensors / scalars)."""
    items = res if isinstance(res, (tuple, list)) else (res,)
    for x in items:
        d = getattr(x, "data", None)
        if d is not None and hasattr(d, "buf"):
            d.buf


def svd_flops(shapes, cx):
    f = 0.0
    for m, n in shapes:
        lo, hi = min(m, n), max(m, n)
        f += 4.0 * hi * lo * lo + 8.0 * lo ** 3
    return f * (4.0 if cx else 1.0)


def qr_flops(shapes, cx):
    f = 0.0
    for m, n in shapes:
        k = min(m, n)
        f += 2.0 * (2.0 * m * n * k - 2.0 * k ** 3 / 3.0)
    return f * (4.0 if cx else 1.0)


def eigh_flops(shapes, cx):
    return sum((4.0 / 3.0 + 8.0) * float(n) ** 3 for n, _ in shapes) * (4.0 if cx else 1.0)


def main():
    args = parse()
    import torch
    from cyten_b200 import tensors as ct
    from cyten_b200.backend import B200AbelianBackend
    from cyten_b200.tensors import SymmetricTensor
    from cyten_b200.workloads import random_blocks

    dev = torch.device("cuda:0")
    be = B200AbelianBackend(dev)
    chi, dt = args.chi, args.dtype
    cx = dt == "complex128"
    esz = 16 if cx else 8
    v, p = legs_for(chi, args.hubbard)
    rows = args.rows.split(",")

    hostA = random_blocks([v, p], [v], 7, dt)
    hostB = random_blocks([v, p], [v], 8, dt)
    hostT = random_blocks([v, p], [p.dual, v], 9, dt)     # full-rank two-site theta
    A = SymmetricTensor.from_numpy_blocks(hostA[1], hostA[0], [v, p], [v], be, labels=["vL", "p", "vR"], dtype=dt)
    B = SymmetricTensor.from_numpy_blocks(hostB[1], hostB[0], [v, p], [v], be, labels=["vL", "p", "vR"], dtype=dt)
    TH = SymmetricTensor.from_numpy_blocks(hostT[1], hostT[0], [v, p], [p.dual, v], be,
                                           labels=["vL", "p0", "p1", "vR"], dtype=dt)
    # Hermitian two-leg tensor on the bond leg for eigh
    rng = np.random.default_rng(11)
    hb, hbi = [], []
    for i, m in enumerate(v.multiplicities):
        x = rng.standard_normal((m, m)) + (1j * rng.standard_normal((m, m)) if cx else 0)
        hb.append(x + x.conj().T)
        hbi.append([i, i])
    HM = SymmetricTensor.from_numpy_blocks(hb, np.asarray(hbi), [v], [v], be, labels=["a", "b"], dtype=dt)

    peak_hbm = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()).get("hbm_gbs", 6650.0) \
        if (ROOT / "MEASURED_PEAKS.json").exists() else 6650.0
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.matmul(a, a)
    best = float("inf")
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, a); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    peak_f64 = 2.0 * n ** 3 / (best * 1e-3) / 1e12
    del a

    THc = ct.combine_legs(TH, [0, 1], [2, 3])
    touch(THc)
    sect = [(int(s[0]), int(s[1])) for s in THc.data.shapes]
    Ac = ct.combine_legs(A, [0, 1])
    qsect = [(int(s[0]), int(s[1])) for s in Ac.data.shapes]
    hsect = [(int(s[0]), int(s[1])) for s in HM.data.shapes]

    out = []

    def emit(row, ms, work, unit, peak, note=""):
        rate = work / (ms * 1e-3) / (1e9 if unit == "GB/s" else 1e12)
        out.append({"row": row, "chi": chi, "dtype": dt, "gpu_ms": ms, "work": work, "rate": rate, "unit": unit,
                    "peak": peak, "frac": rate / peak, "note": note})

    # Kernel-level time of the short rows: one Python call costs more than these kernels run, so besides the eager
    # call (what a user pays per call) 16 calls are captured into ONE CUDA graph and the replay is timed; operands
    # cycle through a pool of 8 clones (larger than L2 at chi=4096), so every call reads from HBM.
    from cyten_b200.data import AbelianBackendData
    NREP = 16

    def clone(T, scale=1.0):
        d = T.data
        buf = d.buf.clone()
        if scale != 1.0:
            buf *= scale
        return SymmetricTensor(AbelianBackendData(d.dtype, d.device, buf, d.block_inds, d.offsets, d.shapes, sig=d._sig),
                               T.codomain, T.domain, be, T.labels)

    def graph_time(body):
        touch(body(0))
        touch(body(1))
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        keep = []
        with torch.cuda.graph(g):
            for i in range(NREP):
                r = body(i)
                touch(r)
                keep.append(r)
        best = float("inf")
        for _ in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        del g, keep
        return best / NREP
    pool = [clone(TH, 1.0 + 0.1 * i) for i in range(8)]

    be.plan_mkn = None
    if "permute" in rows:
        ms = gpu_time(lambda: ct.permute_legs(TH, ["vR", "p1"], ["vL", "p0"]), args.reps, touch)
        emit("a3 permute_legs(theta -> [vR,p1 | vL,p0])", ms, 2.0 * TH.data.nbytes(), "GB/s", peak_hbm)
        out[-1]["graph_ms"] = graph_time(lambda i: ct.permute_legs(pool[i % 8], ["vR", "p1"], ["vL", "p0"]))
    if "combine" in rows:
        ms = gpu_time(lambda: ct.combine_legs(TH, [0, 1], [2, 3]), args.reps, touch)
        emit("a4 combine_legs(theta, [vL,p0], [p1,vR])", ms, 2.0 * TH.data.nbytes(), "GB/s", peak_hbm)
        out[-1]["graph_ms"] = graph_time(lambda i: ct.combine_legs(pool[i % 8], [0, 1], [2, 3]))
    if "split" in rows:
        ms = gpu_time(lambda: ct.split_legs(THc), args.reps, touch)
        emit("a5 split_legs", ms, 2.0 * TH.data.nbytes(), "GB/s", peak_hbm)
        cpool = [clone(THc) for _ in range(8)]
        out[-1]["graph_ms"] = graph_time(lambda i: ct.split_legs(cpool[i % 8]))
        del cpool
    if "tdot" in rows:
        be.plan_mkn = []
        touch(ct.tdot(A, B, [2], [0]))
        fl = (8.0 if cx else 2.0) * float(sum(be.plan_mkn))
        be.plan_mkn = None
        ms = gpu_time(lambda: ct.tdot(A, B, [2], [0]), args.reps, touch)
        emit("a6 tdot(A, B, vR, vL)", ms, fl, "TFLOP/s", peak_f64)
        apool, bpool = [clone(A) for _ in range(8)], [clone(B) for _ in range(8)]
        out[-1]["graph_ms"] = graph_time(lambda i: ct.tdot(apool[i % 8], bpool[i % 8], [2], [0]))
        del apool, bpool
    if "svd" in rows:
        ms = gpu_time(lambda: ct.svd(TH), max(2, args.reps // 2), touch)
        emit("a7 svd(theta) full rank", ms, svd_flops(sect, cx), "TFLOP/s", peak_f64,
             f"{len(sect)} sectors, largest {max(sect)}; sweeps_max={int(be.last_info.abs().max().item())}")
    if "tsvd" in rows:
        ms = gpu_time(lambda: ct.truncated_svd(TH, chi_max=chi, svd_min=1e-14), max(2, args.reps // 2), touch)
        emit("a7+a8 truncated_svd(theta, chi_max=chi)", ms, svd_flops(sect, cx), "TFLOP/s", peak_f64)
    if "qr" in rows:
        ms = gpu_time(lambda: ct.qr(A), max(2, args.reps // 2), touch)
        emit("a9 qr(A[vL,p | vR])", ms, qr_flops(qsect, cx), "TFLOP/s", peak_f64,
             f"{len(qsect)} sectors, largest {max(qsect)}")
    if "eigh" in rows:
        ms = gpu_time(lambda: ct.eigh(HM), max(2, args.reps // 2), touch)
        emit("a10 eigh(H[v | v])", ms, eigh_flops(hsect, cx), "TFLOP/s", peak_f64,
             f"{len(hsect)} sectors, largest {max(hsect)}")

    if "blas1" in rows:
        # (f1) Krylov vector operations.  The kernels take 10-30 us, less than one Python call: 16 calls are captured
        # into one CUDA graph and the replay is timed; operands cycle through a pool of 8 theta-sized arenas (8 x 45 MB
        # at chi=4096 > L2), so every call reads from HBM.
        nb = float(TH.data.nbytes())
        scal = torch.zeros(4, dtype=torch.float64, device=dev)
        ms = graph_time(lambda i: be.inner_dev(pool[(2 * i) % 8].data, pool[(2 * i + 1) % 8].data, True, out=scal[0:2]))
        emit("f1-inner inner(theta, theta')", ms, 2 * nb, "GB/s", peak_hbm)
        ms = graph_time(lambda i: be.norm_sq_dev(pool[i % 8].data, out=scal[0:2]))
        emit("f1-norm norm(theta)", ms, nb, "GB/s", peak_hbm)
        ms = graph_time(lambda i: be.linear_combination(0.3, pool[(3 * i) % 8].data, -1.7, pool[(3 * i + 1) % 8].data))
        emit("f1-lincomb linear_combination(a, theta, b, theta')", ms, 3 * nb, "GB/s", peak_hbm)
        ms = graph_time(lambda i: be.lincomb3_dev([pool[(4 * i) % 8].data, pool[(4 * i + 1) % 8].data, pool[(4 * i + 2) % 8].data],
                                                  [1.0, -1.0, -0.5], coef_dev=[None, scal[0:2], None], sumsq=scal[2:3],
                                                  out=pool[(4 * i + 3) % 8].data))
        emit("f1-lanczos3 w - alpha v - beta v' and its norm (fused, alpha on device)", ms, 4 * nb, "GB/s", peak_hbm)
        u_, s_, vh_ = ct.svd(TH)
        touch(u_)
        upool = [clone(u_, 1.0) for _ in range(8)]
        ms = graph_time(lambda i: be.scale_axis(upool[i % 8].data, s_.data, 2))
        emit("f1-scale scale_axis(U, S, -1)", ms, 2.0 * u_.data.nbytes(), "GB/s", peak_hbm)
        if cx:
            ms = graph_time(lambda i: be.dagger(pool[i % 8].data))
            emit("f1-dagger dagger(theta)", ms, 2 * nb, "GB/s", peak_hbm)

    # ---- the reference on the host cores, same inputs ----
    ref = {}
    if not args.no_ref and list((ROOT / "oracle" / "_ref" / "cyten").glob("_core*.so")):
        sys.path.insert(0, str(ROOT / "oracle" / "_ref"))
        import cyten as rc
        rbe = rc.get_backend("abelian", "numpy")
        sym = rc.u1_symmetry * rc.u1_symmetry if args.hubbard else rc.u1_symmetry
        rdt = rc.Dtype.complex128 if cx else rc.Dtype.float64

        def conv(leg):
            return rc.ElementarySpace.from_defining_sectors(sym, leg.defining_sectors, leg.multiplicities, is_dual=leg.is_dual)

        def mk(host, cod, dom, labels):
            bi, blocks = host
            data = rc.backends.AbelianBackendData(dtype=rdt, device="cpu",
                                                  blocks=[rbe.block_backend.as_block(b) for b in blocks],
                                                  block_inds=rc.backends.BlockInds(bi), is_sorted=True)
            return rc.SymmetricTensor(data, [conv(l) for l in cod], [conv(l) for l in dom], backend=rbe, labels=labels)
        rA = mk(hostA, [v, p], [v], ["vL", "p", "vR"])
        rB = mk(hostB, [v, p], [v], ["vL", "p", "vR"])
        rT = mk(hostT, [v, p], [p.dual, v], ["vL", "p0", "p1", "vR"])
        rH = mk((np.asarray(hbi), hb), [v], [v], ["a", "b"])
        rTc = rc.combine_legs(rT, [0, 1], [2, 3])

        def cpu_time(fn, reps=2):
            fn()
            best = float("inf")
            for _ in range(reps):
                t0 = time.perf_counter()
                fn()
                best = min(best, time.perf_counter() - t0)
            return best * 1e3

        def force(t):   # numpy permutes are lazy views: force like the next consumer would
            for b in t.data.blocks:
                np.ascontiguousarray(rbe.block_backend.to_numpy(b))
        calls = {
            "a3": lambda: force(rc.permute_legs(rT, ["vR", "p1"], ["vL", "p0"], bend_right=None)),
            "a4": lambda: rc.combine_legs(rT, [0, 1], [2, 3]),
            "a5": lambda: rc.split_legs(rTc),
            "a6": lambda: rc.tdot(rA, rB, ["vR"], ["vL"]),
            "a7": lambda: rc.svd(rT),
            "a7+a8": lambda: rc.truncated_svd(rT, chi_max=chi, svd_min=1e-14),
            "a9": lambda: rc.qr(rA),
            "a10": lambda: rc.eigh(rH, ["a", "b", "c"], False),
        }
        if "blas1" in rows:
            rT2 = 1.1 * rT
            rT3 = 1.2 * rT
            rU, rS, rVh = rc.svd(rT)
            lin = rc.tensors.linear_combination
            calls.update({
                "f1-inner": lambda: rc.inner(rT, rT2),
                "f1-norm": lambda: rc.norm(rT),
                "f1-lincomb": lambda: lin(0.3, rT, -1.7, rT2),
                "f1-lanczos3": lambda: rc.norm(lin(1.0, lin(1.0, rT, -0.7, rT2), -0.5, rT3)),
                "f1-scale": lambda: rc.scale_axis(rU, rS, 2),
                "f1-dagger": lambda: force(rc.dagger(rT)),
            })
        for o in out:
            key = o["row"].split()[0]
            try:
                ref[key] = cpu_time(calls[key], 1 if key in ("a7", "a7+a8", "a9", "a10") else 2)
            except Exception as e:   # noqa: BLE001
                ref[key] = f"failed: {e!r}"
    # ---- the reference's GPU path on the same box (comparator only, never the product) ----
    ref_t, ref_t_kind = {}, None
    if args.ref_torch and not args.no_ref and list((ROOT / "oracle" / "_ref" / "cyten").glob("_core*.so")):
        def dev_time(fn, reps=2):
            fn()
            torch.cuda.synchronize()
            best = float("inf")
            for _ in range(reps):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                fn()
                torch.cuda.synchronize()
                best = min(best, time.perf_counter() - t0)
            return best * 1e3
        try:
            tbe = rc.get_backend("abelian", "torch")

            def mv(T):
                T2 = T.to_backend(tbe)
                T2.move_to_device("cuda")
                return T2
            tA, tB, tT, tH = mv(rA), mv(rB), mv(rT), mv(rH)
            tTc = rc.combine_legs(tT, [0, 1], [2, 3])
            tcalls = {
                "a3": lambda: rc.permute_legs(tT, ["vR", "p1"], ["vL", "p0"], bend_right=None),
                "a4": lambda: rc.combine_legs(tT, [0, 1], [2, 3]),
                "a5": lambda: rc.split_legs(tTc),
                "a6": lambda: rc.tdot(tA, tB, ["vR"], ["vL"]),
                "a7": lambda: rc.svd(tT),
                "a7+a8": lambda: rc.truncated_svd(tT, chi_max=chi, svd_min=1e-14),
                "a9": lambda: rc.qr(tA),
                "a10": lambda: rc.eigh(tH, ["a", "b", "c"], False),
            }
            ref_t_kind = "reference AbelianBackend+TorchBlockBackend on cuda"
            for o in out:
                key = o["row"].split()[0]
                if key in tcalls:
                    try:
                        ref_t[key] = dev_time(tcalls[key], 1 if key in ("a7", "a7+a8") else 2)
                    except Exception as e:   # noqa: BLE001
                        ref_t[key] = f"failed: {e!r}"
        except Exception as e:   # noqa: BLE001
            ref_t_kind = f"reference torch backend unusable on cuda ({e!r}); per-sector torch.linalg loop instead"
        # the same library calls without the reference's host layer: torch.linalg.* per sector, torch.matmul per sector
        tdt = torch.complex128 if cx else torch.float64
        smats = [torch.randn(m, n, dtype=tdt, device=dev) for m, n in sect]
        qmats = [torch.randn(m, n, dtype=tdt, device=dev) for m, n in qsect]
        hmats = [torch.as_tensor(b, device=dev) for b in hb]
        loops = {
            "a7": lambda: [torch.linalg.svd(x, full_matrices=False) for x in smats],
            "a7-gesvdj": lambda: [torch.linalg.svd(x, full_matrices=False, driver="gesvdj") for x in smats],
            "a7-gesvda": lambda: [torch.linalg.svd(x, full_matrices=False, driver="gesvda") for x in smats],
            "a9": lambda: [torch.linalg.qr(x, mode="reduced") for x in qmats],
            "a10": lambda: [torch.linalg.eigh(x) for x in hmats],
        }
        loop_ms = {}
        for k, fn in loops.items():
            try:
                loop_ms[k] = dev_time(fn, 2)
            except Exception as e:   # noqa: BLE001
                loop_ms[k] = f"failed: {e!r}"
        print(json.dumps({"row": "cusolver per-sector loops (torch.linalg, same sector shapes)", "chi": chi, "dtype": dt,
                          "ms": loop_ms, "reference_torch_backend": ref_t_kind}))
        for o in out:
            key = o["row"].split()[0]
            o["cusolver_loop_ms"] = loop_ms.get(key)
    for o in out:
        key = o["row"].split()[0]
        o["ref_cpu_ms"] = ref.get(key)
        if ref_t:
            o["ref_torch_cuda_ms"] = ref_t.get(key)
        o["cores"] = os.cpu_count()
        if isinstance(o["ref_cpu_ms"], float):
            o["speedup"] = o["ref_cpu_ms"] / o["gpu_ms"]
        if "graph_ms" in o:
            o["graph_rate"] = o["work"] / (o["graph_ms"] * 1e-3) / (1e9 if o["unit"] == "GB/s" else 1e12)
            o["graph_frac"] = o["graph_rate"] / o["peak"]
        print(json.dumps(o))
    print(f"\n| row | GPU ms | rate | frac of peak | in a graph: ms, rate, frac | reference CPU ms ({os.cpu_count()} cores) | note |", file=sys.stderr)
    print("|---|---|---|---|---|---|---|", file=sys.stderr)
    for o in out:
        r = o["ref_cpu_ms"]
        g = f"{o['graph_ms']:.3f}, {o['graph_rate']:.1f} {o['unit']}, {o['graph_frac']:.3f}" if "graph_ms" in o else ""
        print(f"| {o['row']} | {o['gpu_ms']:.3f} | {o['rate']:.1f} {o['unit']} | {o['frac']:.3f} | {g} | "
              f"{r if not isinstance(r, float) else round(r, 1)} | {o['note']} |", file=sys.stderr)
    if args.out:
        Path(args.out).write_text("\n".join(json.dumps(o) for o in out) + "\n")


if __name__ == "__main__":
    main()

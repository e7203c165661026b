#!/usr/bin/env python3
"""Times the Lanczos loop around the bench's H_eff matvec on one B200.

  python tools/lanczos_probe.py [--chi 4096] [--iters 10] [--ref] [--out gpurun_out/lanczos.json]

N_max = N_min = iters with P_tol = 0 forces exactly ``iters`` iterations.  The environments of the bench H_eff are
random (not Hermitian), so the Ritz values mean nothing: this is a TIMING of matvec + vector operations + the one
read-back per iteration; correctness of the loop is tests/test_krylov.py and tests/test_live_reference.py.
``--ref`` runs cyten.krylov_based.lanczos of the unmodified reference on the host cores with the same operator.
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chi", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from cyten_b200.backend import B200AbelianBackend
    from cyten_b200.krylov import GraphedOperator, LanczosGroundState
    from cyten_b200.workloads import build_heff
    be = B200AbelianBackend("cuda:0")
    H, theta, _, _ = build_heff(args.chi, be)
    opts = {"N_max": args.iters, "N_min": args.iters, "P_tol": 0.0, "N_cache": args.iters}
    res = {"chi": args.chi, "iters": args.iters}

    def timed(fn, reps=3):
        fn()
        best = float("inf")
        for _ in range(reps):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3))
        return best, out
    for _ in range(3):
        H.matvec(theta).data.buf
    ms, _ = timed(lambda: [H.matvec(theta).data.buf for _ in range(args.iters)])
    res["matvec_only_eager_ms_per_iter"] = ms / args.iters
    G = GraphedOperator(H)
    G.matvec(theta)
    ms, _ = timed(lambda: [G.matvec(theta) for _ in range(args.iters)])
    res["matvec_only_graph_ms_per_iter"] = ms / args.iters
    for name, op in (("eager", H), ("graph", G)):
        launches = be.launches
        ms, (E0, psi, N) = timed(lambda: LanczosGroundState(op, theta, dict(opts)).run())
        lz = LanczosGroundState(op, theta, dict(opts))
        lz.run()
        res[f"lanczos_{name}_ms"] = ms
        res[f"lanczos_{name}_ms_per_iter"] = ms / N
        res[f"lanczos_{name}_host_syncs"] = lz.host_syncs
        res["N"] = N
    res["vector_op_overhead_per_iter_ms"] = res["lanczos_graph_ms_per_iter"] - res["matvec_only_graph_ms_per_iter"]
    if args.ref:
        sys.path.insert(0, str(ROOT))
        import bench
        Hr, thr, rc = bench.build_reference_heff(args.chi)
        from cyten import krylov_based, sparse

        class RefOp(sparse.LinearOperator):
            def matvec(self, t):
                return Hr.matvec(t)
        t0 = time.perf_counter()
        Er, psir, Nr = krylov_based.lanczos(RefOp(), thr, dict(opts))
        res["reference_cpu_ms"] = (time.perf_counter() - t0) * 1e3
        res["reference_N"] = int(Nr)
        res["reference_cores"] = __import__("os").cpu_count()
        res["E0_rel_diff_vs_reference"] = abs(E0 - float(Er.to_numpy() if hasattr(Er, "to_numpy") else Er)) / abs(E0)
    print(json.dumps(res))
    if args.out:
        Path(args.out).write_text(json.dumps(res) + "\n")


if __name__ == "__main__":
    main()

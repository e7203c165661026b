"""Worker of tests/test_sharding.py: one process per (virtual) GPU.  ``mode="gloo"``: gloo backend, the host logic of
cyten_b200 over the numpy oracle of the C-ABI (CPU suite).  ``mode="nccl"``: one rank per real GPU, NCCL, the CUDA
library -- the sharded results are compared with the UNSHARDED result computed on the same GPU."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _dense_of(T):
    return T.to_dense()


def worker(rank, world, port, out_dir, mode="gloo", chi=64):
    import torch
    import torch.distributed as dist
    from cyten_b200 import sharding as sh
    from cyten_b200 import tensors as ct
    from cyten_b200.backend import B200AbelianBackend
    from cyten_b200.workloads import CONFIG_LEGS, build_heff, mps_tensor, u1_bond_leg, u1_phys_leg

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if mode == "nccl":
        torch.cuda.set_device(rank)
        dev = torch.device(f"cuda:{rank}")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=world)

    def tt(x, dtype=None):      # scalars for the collectives live where the process group wants them
        return torch.tensor(x, dtype=dtype, device=dev)
    try:
        if mode == "nccl":
            be = B200AbelianBackend(dev)
        else:
            from oracle import cabi_numpy
            be = B200AbelianBackend("cpu", _library=cabi_numpy)
        pg = sh.attach_process_group(be, dist.group.WORLD)
        assert (pg.rank, pg.world) == (rank, world)
        res = {}

        # ---- H_eff matvec: sharded chain + one all_reduce == unsharded ----
        Q, sigma = CONFIG_LEGS.get(chi, (3, 1.2))
        H, theta, _, _ = build_heff(chi, be, Q=Q, sigma=sigma)
        want = H.matvec(theta)
        SH = sh.ShardedHEffective(H, pg)                       # whole vR sectors by flop-cost LPT
        th_loc = SH.shard(theta)
        nblk = tt([th_loc.data.num_blocks])
        dist.all_reduce(nblk)
        assert int(nblk.item()) >= theta.data.num_blocks          # cover (a heavy sector may live on several ranks,
        #                                                            each with a disjoint part of LP's rows)
        be.plan_mkn = []
        got = SH.matvec(th_loc)
        fl = tt([float(sum(be.plan_mkn))], dtype=torch.float64)
        be.plan_mkn = []
        H.matvec(theta)
        fl_full = float(sum(be.plan_mkn))
        be.plan_mkn = None
        dist.all_reduce(fl)
        assert float(fl.item()) == fl_full, "sharding must not duplicate flops"
        full_want = sh.expand_to_full(want)
        assert np.array_equal(got.data.block_inds, full_want.data.block_inds)
        d_got, d_want = _dense_of(got), _dense_of(want)
        assert np.allclose(d_got, d_want, rtol=0, atol=1e-12 * np.abs(d_want).max())
        res["matvec_max"] = float(np.abs(d_want).max())

        # ---- the same with the heavy vR sectors cut into index ranges (slice_leg) ----
        SC = sh.ShardedHEffective(H, pg, split_columns=True, column_align=2 if chi <= 64 else 16, column_heavy=0.25)
        th_c = SC.shard(theta)
        cover = tt([float(th_c.data.sizes.sum())], dtype=torch.float64)
        dist.all_reduce(cover)
        assert float(cover.item()) == float(theta.data.sizes.sum())          # every element of theta on exactly one rank
        be.plan_mkn = []
        got_c = SC.matvec(th_c)
        fl = tt([float(sum(be.plan_mkn))], dtype=torch.float64)
        be.plan_mkn = None
        dist.all_reduce(fl)
        assert float(fl.item()) == fl_full, "the column split must not duplicate flops"
        assert np.array_equal(got_c.data.block_inds, full_want.data.block_inds)
        d_c = _dense_of(got_c)
        assert np.allclose(d_c, d_want, rtol=0, atol=1e-12 * np.abs(d_want).max())
        nsplit = sum(1 for r in SC.ranges for s, (a, b) in r.items() if (a, b) != (0, int(theta.domain.factors[0].multiplicities[s])))
        assert nsplit >= 2, "the test must exercise a sector that is cut into ranges"

        # ---- all_reduce hidden behind the last GEMM: column groups of the result, one collective per group ----
        SO = sh.ShardedHEffective(H, pg, split_columns=True, column_align=2 if chi <= 64 else 16, column_heavy=0.25, fused=True)
        th_o = SO.shard(theta)
        from cyten_b200.fused import FusedHEffective
        HF = FusedHEffective(H)
        HF.matvec(theta)                                   # (the first call also plans the unfused middle: not counted)
        be.plan_mkn = []
        HF.matvec(theta)
        fl_fused = float(sum(be.plan_mkn))                 # the two fat GEMMs of the fused operator, unsharded
        be.plan_mkn = None
        for fractions in ((0.6, 0.4), (0.5, 0.3, 0.2)):
            OV = sh.OverlappedMatvec(SO, th_o, fractions=fractions)
            assert OV.num_parts == len(fractions) and OV.ranges[0][0] == 0
            assert all(OV.ranges[g][1] == OV.ranges[g + 1][0] for g in range(OV.num_parts - 1))      # a partition of the arena
            OV(th_o)
            be.plan_mkn = []
            got_o = OV(th_o)
            fl = tt([float(sum(be.plan_mkn))], dtype=torch.float64)
            be.plan_mkn = None
            dist.all_reduce(fl)
            assert float(fl.item()) == fl_fused, ("the grouped last GEMM must not duplicate flops", float(fl.item()), fl_fused, fl_full, fractions)
            for rep in range(2):                  # the static arena is reused: a second step must not see the first one's sums
                if rep:
                    got_o = OV(th_o)
                assert np.array_equal(got_o.data.block_inds, full_want.data.block_inds)
                d_o = _dense_of(got_o)
                assert np.allclose(d_o, d_want, rtol=0, atol=1e-12 * np.abs(d_want).max()), (fractions, rep)
        if mode == "nccl":                        # the same step as one CUDA graph per part, collectives between the replays
            from cyten_b200.graphs import CapturedPipeline
            cap = CapturedPipeline([lambda t, g=g: OV.part(g, t) for g in range(OV.num_parts)], th_o)
            for rep in range(2):
                d_g = _dense_of(OV(th_o, run_part=cap.replay))
                assert np.allclose(d_g, d_want, rtol=0, atol=1e-12 * np.abs(d_want).max()), ("graphs", rep)
            res["overlap_graph_launches"] = cap.launches

        # ---- the Krylov loop on top of the sharded matvec: same iterations, same Ritz value as single process ----
        from types import SimpleNamespace
        from cyten_b200.krylov import lanczos
        opts = {"N_max": 6, "N_min": 6, "P_tol": 0.0}
        E1, psi1, N1 = lanczos(H, theta, dict(opts))
        for SX in (SC, SH):
            E2, psi2, N2 = lanczos(SimpleNamespace(matvec=SX.matvec_replicated), theta, dict(opts))
            assert N1 == N2 and abs(E1 - E2) <= 1e-10 * (abs(E1) + 1.0)
            full1, full2 = sh.expand_to_full(psi1), sh.expand_to_full(psi2)
            assert np.allclose(_dense_of(full2), _dense_of(full1), rtol=0, atol=1e-9)

        # ---- tdot with the output blocks partitioned over the ranks ----
        v, p = (u1_bond_leg(48, 3, 1.2) if chi <= 64 else u1_bond_leg(chi, Q, sigma)), u1_phys_leg()
        A, B = mps_tensor(v, p, v, be, seed=1), mps_tensor(v, p, v, be, seed=2)
        th = ct.tdot(A, B, [2], [0])
        th_loc = sh.tdot_sharded(A, B, [2], [0], pg)
        nb = tt([th_loc.data.num_blocks])
        dist.all_reduce(nb)
        assert int(nb.item()) == th.data.num_blocks               # every output block computed exactly once
        th_rep = sh.all_gather_blocks(th_loc, pg)
        assert np.array_equal(_dense_of(th_rep), _dense_of(th))   # same kernels, same order: bit-exact

        # ---- truncated SVD with the sectors LPT-partitioned, global cut via all_gather of S ----
        for chi_max in ((30, 48) if chi <= 64 else (chi // 2, chi)):
            U1, S1, V1, err1, ren1 = ct.truncated_svd(th, chi_max=chi_max, svd_min=1e-12)
            U, S, Vh, err, ren, owner = sh.truncated_svd_sharded(th, pg, chi_max=chi_max, svd_min=1e-12)
            assert abs(err - err1) <= 1e-14 and ren == ren1
            assert S.leg == S1.leg                                # identical kept-sector dimensions
            # my blocks are exactly the sectors I own, and equal the unsharded ones
            s_un = {int(r[0]): b for r, b in zip(S1.data.block_inds, S1.data.to_numpy_blocks())}
            for r, b in zip(S.data.block_inds, S.data.to_numpy_blocks()):
                assert np.array_equal(b, s_un[int(r[0])])
            kept = tt([int(S.data.shapes.sum())])
            dist.all_reduce(kept)
            assert int(kept.item()) == int(S1.data.shapes.sum())
            Ur, Vr = sh.all_gather_blocks(U, pg), sh.all_gather_blocks(Vh, pg)
            assert np.array_equal(_dense_of(Ur), _dense_of(sh.expand_to_full(U1)))
            assert np.array_equal(_dense_of(Vr), _dense_of(sh.expand_to_full(V1)))
        # ---- BASELINE config 5: one bond update of a sweep on N ranks == single process ----
        from cyten_b200.workloads import build_sweep_step
        schi = 64 if chi <= 64 else chi
        sQ, ssig = CONFIG_LEGS.get(schi, (3, 1.2))
        st1, th_s = build_sweep_step(schi, be, chi_max=(3 * schi) // 4, Q=sQ, sigma=ssig)
        r1 = st1.update_bond(th_s)
        stN, _ = build_sweep_step(schi, be, chi_max=(3 * schi) // 4, Q=sQ, sigma=ssig, fused=True)
        stN = sh.shard_sweep_step(stN, pg, column_heavy=0.25, column_align=2 if chi <= 64 else 16)
        rN = stN.update_bond(th_s)
        assert abs(rN["err"] - r1["err"]) <= 1e-13
        assert rN["S"].leg == r1["S"].leg
        sN = np.concatenate([x for x in rN["S"].data.to_numpy_blocks()] or [np.zeros(0)])      # replicated already
        s1 = np.concatenate([x for x in r1["S"].data.to_numpy_blocks()] or [np.zeros(0)])
        assert np.allclose(np.sort(sN[sN > 0]), np.sort(s1), rtol=0, atol=1e-12 * s1.max())
        # the sharded and the unsharded matvec differ by rounding, so the singular vectors may differ by a sign per
        # vector: the environments are compared entry by entry in absolute value
        for name in ("LP", "RP", "Bi", "Bj"):
            a, b = _dense_of(sh.expand_to_full(rN[name])), _dense_of(sh.expand_to_full(r1[name]))
            assert np.allclose(np.abs(a), np.abs(b), rtol=0, atol=1e-7 * np.abs(b).max()), name
        res["ok"] = True
        np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.asarray([1.0]))
    finally:
        dist.destroy_process_group()

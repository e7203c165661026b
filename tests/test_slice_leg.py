"""slice_leg / column_ranges: the finer-than-sector unit of the multi-GPU split (cyten_b200/sharding.py), single process."""
from __future__ import annotations

import numpy as np
import pytest

from cyten_b200 import sharding as sh
from cyten_b200 import tensors as ct
from cyten_b200.workloads import HEFF_LABELS, HEffective, build_heff


@pytest.mark.parametrize("dtype", ["float64", "complex128"])
def test_slice_leg_matches_dense_slicing(backend, dtype):
    H, theta, _, _ = build_heff(64, backend, dtype=dtype, Q=3, sigma=1.2)
    leg = theta.domain.factors[0]                       # vR
    m = leg.multiplicities
    ranges = {1: (0, int(m[1])), 3: (2, int(m[3]) - 3), 4: (1, 2)}
    part = sh.slice_leg(theta, "vR", ranges)
    assert part.labels == theta.labels
    new_leg = part.domain.factors[0]
    assert np.array_equal(new_leg.defining_sectors, leg.defining_sectors[[1, 3, 4]])
    assert list(new_leg.multiplicities) == [int(m[1]), int(m[3]) - 5, 1]
    dense = theta.to_dense()
    cols = np.concatenate([np.arange(leg.slices[s][0] + a, leg.slices[s][0] + b) for s, (a, b) in sorted(ranges.items())])
    assert np.array_equal(part.to_dense(), dense[..., cols])       # vR is the last block axis
    # first axis of RP (the partner of theta's vR in the last contraction)
    rp = sh.slice_leg(H.RP, 0, ranges)
    assert np.array_equal(rp.to_dense(), H.RP.to_dense()[cols])
    with pytest.raises(ValueError):
        sh.slice_leg(theta, "vR", {1: (0, int(m[1]) + 1)})


def test_partial_chains_add_up(backend):
    """H_eff(theta) = sum over the ranges of H_eff restricted to (theta_r, RP_r): what the ranks compute."""
    H, theta, _, _ = build_heff(64, backend, Q=3, sigma=1.2)
    want = sh.expand_to_full(H.matvec(theta)).to_dense()
    leg = theta.domain.factors[0]
    costs = sh.flop_costs_of_leg(theta, "vR", H.matvec)
    for world in (2, 3, 5):
        ranges = sh.column_ranges(costs, leg.multiplicities, world, heavy=0.25, align=2)
        covered = np.zeros(leg.dim, dtype=np.int64)
        acc = np.zeros_like(want)
        for r in ranges:
            for s, (a, b) in r.items():
                covered[leg.slices[s][0] + a:leg.slices[s][0] + b] += 1
            if not r:
                continue
            Hr = HEffective(H.LP, sh.slice_leg(H.RP, 0, r), H.W1, H.W2, H.ct)
            acc += sh.expand_to_full(Hr.matvec(sh.slice_leg(theta, "vR", r))).to_dense()
        assert np.all(covered[np.repeat(costs > 0, leg.multiplicities)] == 1)      # every index on exactly one rank
        assert np.allclose(acc, want, rtol=0, atol=1e-12 * np.abs(want).max())
        loads = [sum(costs[s] * (b - a) / leg.multiplicities[s] for s, (a, b) in r.items()) for r in ranges]
        whole = sh.lpt_imbalance(costs, sh.lpt_assign(costs, world), world)
        assert max(loads) / (sum(loads) / world) <= whole + 0.1                     # (LPT is a heuristic) about as good or better


def test_valid_block_inds_join_equals_brute_force():
    """valid_block_inds joins the codomain and domain grids on the fused charge; the result must
    be the reference's brute-force list (abelian.cpp:329-399), lexsorted, for U(1) x U(1) and six legs."""
    from cyten_b200.spaces import TensorProduct
    from cyten_b200.tensors import valid_block_inds
    from cyten_b200.workloads import hubbard_bond_leg, hubbard_phys_leg
    v, p = hubbard_bond_leg(256, Q=3), hubbard_phys_leg()
    cod, dom = TensorProduct([v, p, p]), TensorProduct([v, p.dual, p.dual])
    bi = valid_block_inds(cod, dom)
    legs = list(cod.factors) + list(reversed(dom.factors))
    sym = cod.symmetry
    grid = np.indices([l.num_sectors for l in legs]).reshape(len(legs), -1).T
    qc = sym.multiple_fusion_broadcast([l.sector_decomposition[grid[:, i]] for i, l in enumerate(legs[:3])])
    qd = sym.multiple_fusion_broadcast([l.sector_decomposition[grid[:, 3 + i]] for i, l in enumerate(legs[3:])])
    want = grid[np.all(qc == qd, axis=1)]
    want = want[np.lexsort(want.T)]
    assert len(bi) > 100 and np.array_equal(bi, want)
    assert valid_block_inds(TensorProduct([p]), TensorProduct([v], symmetry=sym)).shape[1] == 2


def test_plan_only_mode_refuses_arena_kernels(backend):
    """``dry_run`` arenas are 16-element dummies: an operation that would launch a kernel on them must raise, not fault."""
    import pytest
    from cyten_b200 import tensors as ct
    from cyten_b200._lib import CytenB200Error
    from cyten_b200.workloads import mps_tensor, u1_bond_leg, u1_phys_leg
    v, p = u1_bond_leg(24, 2, 1.0), u1_phys_leg()
    A = mps_tensor(v, p, v, backend, seed=1)
    backend.dry_run = True
    try:
        th = ct.tdot(A, A, [2], [0])                       # bookkeeping only: fine
        assert th.data.num_blocks > 0
        with pytest.raises(CytenB200Error, match="plan-only"):
            ct.svd(th)
    finally:
        backend.dry_run = False
    assert float(ct.norm(ct.tdot(A, A, [2], [0]))) > 0     # the backend is usable again

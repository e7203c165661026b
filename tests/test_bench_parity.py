"""Value-level parity AT THE BENCHMARK CONFIGURATIONS against the unmodified reference run live on the host cores
(oracle/_ref, built by oracle/build_ref.py; it travels to the GPU box as a built artefact).

* chi = 4096 H_eff matvec of ``bench.py`` (BASELINE config 2 at the size the metric is quoted on): every block of
  ``y`` against the reference's ``y``, ``max|y - y_ref| <= 1e-12 * ||y_ref||``, equal ``block_inds``.
* chi = 4096 ``truncated_svd`` (BASELINE config 3): singular values against the reference's (<= 1e-12 * sigma_max),
  kept sector dimensions equal, ``err`` equal; every sector of the full SVD against LAPACK (numpy) and
  ``U diag(S) Vh = theta`` evaluated with numpy, not with the GPU's own ``compose``.
* BASELINE config 4 at its size (Fermi-Hubbard U(1) x U(1), chi = 8192, complex128): ``tdot`` against numpy on a sample
  of output blocks, SVD of the heaviest sectors against LAPACK.  The full reference is too slow at this size (minutes
  per call on the box's host cores), hence the sample.

What the reference asserts for the same operations: tests/python_tests/test_tensors.py:3393-3510, 3754-3764.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
REF = ROOT / "oracle" / "_ref"
HAVE_REF = bool(list((REF / "cyten").glob("_core*.so")))

from cyten_b200 import tensors as ct  # noqa: E402
from cyten_b200.spaces import TensorProduct  # noqa: E402
from cyten_b200.workloads import (CONFIG_LEGS, build_heff, hubbard_bond_leg, hubbard_phys_leg, mps_tensor,  # noqa: E402
                                  random_blocks, random_tensor, u1_bond_leg, u1_phys_leg)

needs_ref = pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref is not built (python oracle/build_ref.py)")


def _ref_blocks(T):
    bi = np.asarray(T.data.block_inds.to_numpy()).reshape(len(T.data.blocks), -1)
    return bi, [np.asarray(b.to_numpy()) for b in T.data.blocks]


# (backend, chi): the numpy oracle of the C-ABI at a size it finishes in seconds (CPU suite: checks this file's logic and
# the host layer), the CUDA library at the benchmark size
SIZES = [pytest.param("oracle", 128, id="oracle-128"), pytest.param("cuda", 4096, marks=pytest.mark.gpu, id="cuda-4096")]


def _backend(kind):
    from conftest import _cuda_backend, _oracle_backend
    return _cuda_backend() if kind == "cuda" else _oracle_backend()


@needs_ref
@pytest.mark.parametrize("kind,chi", SIZES)
def test_heff_matvec_matches_live_reference(kind, chi):
    import bench
    H, theta, _, _ = build_heff(chi, _backend(kind))
    y = H.matvec(theta)
    Href, theta_ref, _ = bench.build_reference_heff(chi)      # the same host blocks (workloads.random_blocks, same seeds)
    yref = Href.matvec(theta_ref)
    bi_ref, blocks_ref = _ref_blocks(yref)
    assert list(yref.labels) == list(y.labels)
    assert np.array_equal(y.data.block_inds, bi_ref)
    nrm = np.sqrt(sum(float(np.vdot(b, b).real) for b in blocks_ref))
    worst = 0.0
    for g, r in zip(y.data.to_numpy_blocks(), blocks_ref):
        assert g.shape == r.shape
        worst = max(worst, float(np.abs(g - r).max()))
    assert worst <= 1e-12 * nrm, (worst, nrm)


def _sector_matrices(theta):
    """theta[vL,p0 | p1,vR] -> the 2-leg tensor whose blocks are the coupled-sector matrices (combine_legs is pure data
    movement, bit-exact against the reference in tests/test_live_reference.py and tests/test_golden_ops.py)."""
    tens, new_co_domain, _, _ = ct._decomposition_prepare(theta, False)
    return tens, new_co_domain


@needs_ref
@pytest.mark.parametrize("kind,chi", SIZES)
def test_truncated_svd_matches_live_reference(kind, chi):
    import bench
    from cyten_b200.interop import diagonal_to_numpy
    be = _backend(kind)
    Q, sigma = CONFIG_LEGS[chi]
    v, p = u1_bond_leg(chi, Q, sigma), u1_phys_leg()
    theta = random_tensor([v, p], [p.dual, v], be, seed=9, labels=["vL", "p0", "p1", "vR"])
    theta_ref, rc = bench.build_reference_theta(chi)
    # identical inputs on both sides
    bi_ref, blocks_ref = _ref_blocks(theta_ref)
    assert np.array_equal(theta.data.block_inds, bi_ref)
    assert all(np.array_equal(a, b) for a, b in zip(theta.data.to_numpy_blocks(), blocks_ref))
    Ur, Sr, Vr, err_r, ren_r = rc.truncated_svd(theta_ref, chi_max=chi, svd_min=1e-14)
    U, S, Vh, err, ren = ct.truncated_svd(theta, chi_max=chi, svd_min=1e-14)
    assert np.array_equal(np.asarray(Sr.leg.multiplicities), S.leg.multiplicities)          # kept sector dimensions
    assert np.array_equal(np.asarray(Sr.leg.defining_sectors).reshape(S.leg.defining_sectors.shape), S.leg.defining_sectors)
    s_ref, s_our = diagonal_to_numpy(Sr), S.to_numpy()
    assert np.abs(s_our - s_ref).max() <= 1e-12 * s_ref.max()
    assert abs(err - float(err_r)) <= 1e-12 and abs(ren - float(ren_r)) <= 1e-12
    # block_inds of the factors equal the reference's
    for ours, ref in ((U, Ur), (Vh, Vr)):
        assert np.array_equal(ours.data.block_inds, _ref_blocks(ref)[0])
    # truncated reconstruction error, evaluated on the host with numpy: ||theta - U S Vh||^2 = err * ||theta||^2
    nrm2 = sum(float(np.vdot(b, b).real) for b in blocks_ref)
    Uc, Vc = ct.combine_legs(U, [0, 1]), ct.combine_legs(Vh, [1, 2])
    tens, _ = _sector_matrices(theta)
    sb = {int(r[0]): b for r, b in zip(S.data.block_inds, S.data.to_numpy_blocks())}
    diff2 = 0.0
    u_by_small = {int(r[1]): (int(r[0]), b) for r, b in zip(Uc.data.block_inds, Uc.data.to_numpy_blocks())}
    v_by_small = {int(r[0]): (int(r[1]), b) for r, b in zip(Vc.data.block_inds, Vc.data.to_numpy_blocks())}
    th_blocks = {(int(r[0]), int(r[1])): b for r, b in zip(tens.data.block_inds, tens.data.to_numpy_blocks())}
    seen = set()
    for k, s in sb.items():
        (j, ublk), (kk, vblk) = u_by_small[k], v_by_small[k]
        rec = (ublk * s[None, :]) @ vblk
        diff2 += float(np.linalg.norm(th_blocks[(j, kk)] - rec) ** 2)
        seen.add((j, kk))
    for key, blk in th_blocks.items():
        if key not in seen:
            diff2 += float(np.linalg.norm(blk) ** 2)
    assert abs(diff2 / nrm2 - err) <= 1e-11, (diff2 / nrm2, err)


@pytest.mark.parametrize("kind,chi", SIZES)
def test_full_svd_sectors_vs_lapack(kind, chi):
    """Every coupled sector of the two-site theta: S against LAPACK (numpy), U diag(S) Vh = block and the isometries
    evaluated with numpy on the host."""
    be = _backend(kind)
    Q, sigma = CONFIG_LEGS[chi]
    v, p = u1_bond_leg(chi, Q, sigma), u1_phys_leg()
    theta = random_tensor([v, p], [p.dual, v], be, seed=9, labels=["vL", "p0", "p1", "vR"])
    tens, new_co_domain = _sector_matrices(theta)
    u, s, vh = be.svd(tens, new_co_domain)
    _check_sector_svd(tens, u, s, vh, 1e-12)


def _check_sector_svd(tens, u, s, vh, tol, sample=None):
    ub = {int(r[1]): (int(r[0]), b) for r, b in zip(u.block_inds, u.to_numpy_blocks())}
    vb = {int(r[0]): (int(r[1]), b) for r, b in zip(vh.block_inds, vh.to_numpy_blocks())}
    th = {(int(r[0]), int(r[1])): b for r, b in zip(tens.data.block_inds, tens.data.to_numpy_blocks())}
    items = list(zip(s.block_inds, s.to_numpy_blocks()))
    if sample is not None:
        items = sorted(items, key=lambda it: -len(it[1]))[:sample]
    assert items
    for r, sv in items:
        n = int(r[0])
        (j, ublk), (k, vblk) = ub[n], vb[n]
        blk = th[(j, k)]
        s_lapack = np.linalg.svd(blk, compute_uv=False)
        smax = max(float(s_lapack.max()), 1e-300)
        assert np.all(np.diff(sv) <= 0) and np.all(sv >= 0)
        assert np.abs(sv - s_lapack).max() <= tol * smax, (n, blk.shape)
        assert np.abs((ublk * sv[None, :]) @ vblk - blk).max() <= 10 * tol * smax * np.sqrt(max(blk.shape)), (n, blk.shape)
        kdim = len(sv)
        assert np.abs(ublk.conj().T @ ublk - np.eye(kdim)).max() <= 1e-11
        assert np.abs(vblk @ vblk.conj().T - np.eye(kdim)).max() <= 1e-11


@pytest.mark.parametrize("kind,chi", [pytest.param("oracle", 256, id="oracle-256"),
                                      pytest.param("cuda", 8192, marks=pytest.mark.gpu, id="cuda-8192")])
def test_config4_hubbard_complex128_sampled(kind, chi):
    """BASELINE config 4 at its stated size: U(1) x U(1) (charge, 2Sz), chi = 8192, d = 4, complex128."""
    be = _backend(kind)
    v, p = hubbard_bond_leg(chi), hubbard_phys_leg()
    bia, blocks_a = random_blocks([v, p], [v], 1, "complex128")
    bib, blocks_b = random_blocks([v, p], [v], 2, "complex128")
    A = ct.SymmetricTensor.from_numpy_blocks(blocks_a, bia, [v, p], [v], be, labels=["vL", "p", "vR"], dtype="complex128")
    B = ct.SymmetricTensor.from_numpy_blocks(blocks_b, bib, [v, p], [v], be, labels=["vL", "p", "vR"], dtype="complex128")
    th = ct.tdot(A, B, [2], [0])                       # [vL, p0 | p1, vR]
    assert th.data.num_blocks > (1000 if chi >= 4096 else 100)      # "many small irregular blocks"
    # numpy on a sample of output blocks: the largest, the smallest and a spread in between
    a_by = {}
    for r, blk in zip(bia, blocks_a):
        a_by.setdefault((int(r[0]), int(r[1])), []).append((int(r[2]), blk))
    b_by = {}
    for r, blk in zip(bib, blocks_b):                  # B blocks have axes [vL, p, vR]
        b_by.setdefault((int(r[1]), int(r[2])), []).append((int(r[0]), blk))
    sizes = th.data.sizes
    order = np.argsort(sizes)
    pick = sorted(set(order[-6:].tolist() + order[:4].tolist() + order[::max(1, len(order) // 30)].tolist()))
    host = th.data.buf.cpu().numpy()
    scale = float(ct.norm(th)) / np.sqrt(th.data.num_blocks)
    for n in pick:
        i, j, k, l = (int(x) for x in th.data.block_inds[n])      # axes [vL, p0, p1, vR]; domain = [vR, p1] reversed
        shape = tuple(int(x) for x in th.data.shapes[n])
        got = host[th.data.offsets[n]:th.data.offsets[n] + sizes[n]].reshape(shape)
        want = np.zeros(shape, dtype=np.complex128)
        bmap = dict(b_by.get((k, l), []))
        for c, ablk in a_by.get((i, j), []):
            if c in bmap:
                want += np.tensordot(ablk, bmap[c], axes=([2], [0]))
        assert np.abs(got - want).max() <= 1e-12 * max(scale, float(np.abs(want).max())) * np.sqrt(chi), (n, shape)
    del host, blocks_a, blocks_b
    # SVD of a full-rank theta at this size: the heaviest sectors against LAPACK (each of them is a ~2000 x 2000 complex matrix)
    theta = random_tensor([v, p], [p.dual, v], be, seed=3, dtype="complex128", labels=["vL", "p0", "p1", "vR"])
    tens, new_co_domain = _sector_matrices(theta)
    u, s, vh = be.svd(tens, new_co_domain)
    assert be.last_info is not None and int(be.last_info.min().item()) > 0          # every sector converged
    _check_sector_svd(tens, u, s, vh, 1e-12, sample=3)
    nrm = float(ct.norm(theta))
    tot = float(sum(np.sum(b ** 2) for b in s.to_numpy_blocks()))
    assert abs(tot - nrm ** 2) <= 1e-12 * nrm ** 2

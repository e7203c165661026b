"""BASELINE config 5: one bond update of a two-site DMRG sweep (toycodes/tenpy_toycodes/d_dmrg.py:217-274) -- the SAME call
sequence (cyten_b200.workloads.SweepStep) on the unmodified reference and on the B200 path, from identical host blocks.

Compared: the new Schmidt values and the truncation error (gauge invariant, 1e-12), the kept sector dimensions (equal),
and the new environments LP / RP entry by entry in absolute value -- the singular vectors of the two SVDs differ by a
sign per vector, which multiplies LP[a, w, b] by s_a s_b (tolerance 1e-9 of the largest entry: the environments
contain 1/S factors of the synthetic left Schmidt values)."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
REF = ROOT / "oracle" / "_ref"
if not list((REF / "cyten").glob("_core*.so")):
    pytest.skip("oracle/_ref is not built (python oracle/build_ref.py)", allow_module_level=True)
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402

from cyten_b200.workloads import build_sweep_step  # noqa: E402


def _ref_blocks(T):
    bi = np.asarray(T.data.block_inds.to_numpy())
    return bi, [np.asarray(b.to_numpy()) for b in T.data.blocks]


def _check(chi, chi_max, backend, fused=False):
    st, th = build_sweep_step(chi, backend, chi_max=chi_max, fused=fused)
    r = st.update_bond(th)
    st0, th0, ct = bench.build_reference_sweep_step(chi, chi_max=chi_max)
    r0 = st0.update_bond(th0)
    assert abs(r["err"] - r0["err"]) < 1e-12
    # kept dimensions per sector and the Schmidt values
    bi0, s0 = _ref_blocks(r0["S"])
    assert np.array_equal(r["S"].data.block_inds, bi0.reshape(r["S"].data.block_inds.shape))
    s = r["S"].data.to_numpy_blocks()
    assert [len(x) for x in s] == [len(x) for x in s0]
    smax = max(float(np.max(x)) for x in s0)
    for x, y in zip(s, s0):
        assert np.allclose(x, y, rtol=0, atol=1e-12 * smax)
    for name in ("LP", "RP", "Bi", "Bj"):
        bi0, blocks0 = _ref_blocks(r0[name])
        assert np.array_equal(r[name].data.block_inds, bi0), name
        blocks = r[name].data.to_numpy_blocks()
        big = max(float(np.abs(b).max()) for b in blocks0)
        for x, y in zip(blocks, blocks0):
            assert x.shape == y.shape, name
            assert np.allclose(np.abs(x), np.abs(y), rtol=0, atol=1e-9 * big), name
    return st


@pytest.mark.parametrize("fused", [False, True])
def test_sweep_step_matches_reference_oracle(oracle_backend, fused):
    _check(128, 100, oracle_backend, fused)


@pytest.mark.gpu
@pytest.mark.parametrize("chi,chi_max", [(128, 100), (1024, 800)])
def test_sweep_step_matches_reference_cuda(cuda_backend, chi, chi_max):
    _check(chi, chi_max, cuda_backend, fused=(chi > 128))

"""The drop-in boundary: ``integration/cyten_plugin.py`` behind tenpy/cyten's own Python API.

* ``cyten.tdot`` / ``truncated_svd`` / ``qr`` / ``lq`` / ``eigh`` / ``permute_legs`` / ``combine_legs`` / ``split_legs`` and the
  Krylov vector operations are called on tensors that live on the plugin backend; the results are compared with the same
  calls on the reference's ``AbelianBackend + NumpyBlockBackend``, and the plugin's counters must show that every hot
  virtual ran on the kernels and none fell through to the inner reference backend.
* The reference's OWN pytest subset (tests/python_tests/test_tensors.py, the ``AbelianBackend`` ids of
  tdot / compose / permute_legs / combine_split / svd / qr_lq / eigh) is run through the plugin in a subprocess
  (``-p pytest_cyten_b200 --block-backends torch``).

CPU leg: the plugin's host logic over the numpy restatement of the C-ABI (``oracle/cabi_numpy``, test infrastructure).
GPU leg (``-m gpu``): the CUDA library.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
REF = ROOT / "oracle" / "_ref"
if not list((REF / "cyten").glob("_core*.so")):
    pytest.skip("oracle/_ref is not built (python oracle/build_ref.py)", allow_module_level=True)
for p in (str(REF), str(ROOT / "integration")):
    if p not in sys.path:
        sys.path.insert(0, p)
import cyten as rc  # noqa: E402

from cyten_plugin import CytenB200Backend, install, to_plugin, to_reference  # noqa: E402

REF_SUBSET = "AbelianBackend and (tdot or compose or permute_legs or combine_split or svd or qr_lq or eigh)"


@pytest.fixture(params=[pytest.param("oracle"), pytest.param("cuda", marks=pytest.mark.gpu)])
def plugin(request) -> CytenB200Backend:
    if request.param == "oracle":
        from oracle import cabi_numpy
        be = install("cpu", key="b200-test-cpu", _library=cabi_numpy)
    else:
        import torch
        if not torch.cuda.is_available():
            pytest.skip("no CUDA device")
        be = install("cuda:0", key="b200-test-gpu")
    for c in (be.hot_calls, be.cold_calls, be.fallbacks):
        c.clear()
    return be


def _legs(sym_name):
    if sym_name == "u1":
        sym = rc.u1_symmetry
        v = rc.ElementarySpace.from_defining_sectors(sym, [[-2], [-1], [0], [1], [2]], [3, 7, 11, 6, 2])
        p = rc.ElementarySpace.from_defining_sectors(sym, [[-1], [1]], [1, 1])
    else:
        sym = rc.u1_symmetry * rc.z3_symmetry
        v = rc.ElementarySpace.from_defining_sectors(sym, [[-1, 0], [0, 1], [0, 2], [1, 1], [2, 0]], [4, 5, 3, 6, 2])
        p = rc.ElementarySpace.from_defining_sectors(sym, [[-1, 1], [1, 2]], [1, 2])
    return sym, v, p


def _pair(sym_name, dtype, plugin, seed=3):
    sym, v, p = _legs(sym_name)
    np.random.seed(seed)
    ref = rc.get_backend("abelian", "numpy")
    rdt = rc.Dtype.complex128 if dtype == "complex128" else rc.Dtype.float64
    A0 = rc.SymmetricTensor.from_random_normal([v, p], [v], backend=ref, labels=["vL", "p", "vR"], dtype=rdt)
    B0 = rc.SymmetricTensor.from_random_normal([v, p], [v], backend=ref, labels=["vL", "p", "vR"], dtype=rdt)
    return A0, B0, to_plugin(A0, plugin), to_plugin(B0, plugin)


def _close(T_plugin, T_ref, tol=1e-12):
    T = to_reference(T_plugin)
    assert np.array_equal(np.asarray(T.data.block_inds.to_numpy()), np.asarray(T_ref.data.block_inds.to_numpy()))
    for x, y in zip(T.data.blocks, T_ref.data.blocks):
        x, y = np.asarray(x.to_numpy()), np.asarray(y.to_numpy())
        assert x.shape == y.shape
        assert np.allclose(x, y, rtol=0, atol=tol * max(1.0, float(np.abs(y).max()) if y.size else 1.0))


CASES = [("u1", "float64"), ("u1", "complex128"), ("u1z3", "float64")]


def test_install_registers_the_backend(plugin):
    key = [k for k, v in rc._core._tensor_backend_cache.items() if v is plugin]
    assert key and key[0][0] == "abelian"
    assert rc.get_backend(*key[0]) is plugin
    assert isinstance(plugin, rc.backends.TensorBackend)
    assert set(plugin.HOT) <= set(dir(plugin))


@pytest.mark.parametrize("sym_name,dtype", CASES)
def test_tdot_through_cyten_api(plugin, sym_name, dtype):
    A0, B0, A, B = _pair(sym_name, dtype, plugin)
    assert A.backend is plugin
    th0 = rc.tdot(A0, B0, ["vR"], ["vL"])
    th = rc.tdot(A, B, ["vR"], ["vL"])
    assert th.backend is plugin
    _close(th, th0, 1e-13)
    assert plugin.hot_calls["compose"] == 1 and plugin.hot_calls["permute_legs"] >= 1
    assert not plugin.fallbacks
    assert plugin.fast.launches > 0          # C-ABI calls that launch cyb:: kernels


@pytest.mark.parametrize("sym_name,dtype", CASES)
def test_leg_manipulation_is_bit_exact(plugin, sym_name, dtype):
    A0, B0, A, B = _pair(sym_name, dtype, plugin)
    th0 = rc.tdot(A0, B0, ["vR"], ["vL"])
    th = to_plugin(th0, plugin)              # the same bits on both sides: pure data movement must reproduce them
    _close(rc.permute_legs(th, [3, 1], [0, 2]), rc.permute_legs(th0, [3, 1], [0, 2]), 0.0)
    c0, c = rc.combine_legs(th0, [0, 1], [2, 3]), rc.combine_legs(th, [0, 1], [2, 3])
    _close(c, c0, 0.0)
    _close(rc.split_legs(c), rc.split_legs(c0), 0.0)
    assert not plugin.fallbacks
    assert plugin.hot_calls["combine_legs"] == 1 and plugin.hot_calls["split_legs"] == 1


@pytest.mark.parametrize("sym_name,dtype", CASES)
def test_truncated_svd_through_cyten_api(plugin, sym_name, dtype):
    A0, B0, A, B = _pair(sym_name, dtype, plugin)
    th0, th = rc.tdot(A0, B0, ["vR"], ["vL"]), rc.tdot(A, B, ["vR"], ["vL"])
    U0, S0, Vh0, err0, ren0 = rc.truncated_svd(th0, chi_max=17, svd_min=1e-12)
    U, S, Vh, err, ren = rc.truncated_svd(th, chi_max=17, svd_min=1e-12)
    # kept sector dimensions and the singular values
    assert np.array_equal(np.asarray(S.leg.multiplicities), np.asarray(S0.leg.multiplicities))
    _close(S, S0, 1e-12)
    assert abs(err - err0) < 1e-12 and abs(ren - ren0) < 1e-12
    # U S Vh is the same truncated tensor (the factors themselves are gauge dependent)
    rec = to_reference(rc.compose(U, rc.compose(S, Vh)))
    rec0 = rc.compose(U0, rc.compose(S0, Vh0))
    assert rc.almost_equal(rec, rec0, rtol=1e-10, atol=1e-10)
    for name in ("combine_legs", "svd", "split_legs", "truncate_singular_values", "mask_contract_large_leg",
                 "apply_mask_to_DiagonalTensor", "norm"):
        assert plugin.hot_calls[name] >= 1, name
    assert not plugin.fallbacks


@pytest.mark.parametrize("sym_name,dtype", CASES)
def test_qr_lq_eigh_through_cyten_api(plugin, sym_name, dtype):
    A0, B0, A, B = _pair(sym_name, dtype, plugin)
    th0, th = rc.tdot(A0, B0, ["vR"], ["vL"]), rc.tdot(A, B, ["vR"], ["vL"])
    Q, R = rc.qr(th)
    assert rc.almost_equal(to_reference(rc.compose(Q, R)), th0, rtol=1e-11, atol=1e-11)
    L, Q2 = rc.lq(th)
    assert rc.almost_equal(to_reference(rc.compose(L, Q2)), th0, rtol=1e-11, atol=1e-11)
    H0 = rc.compose(th0, th0.hc)
    H = rc.compose(th, th.hc)
    _close(H, H0, 1e-12)
    for sort in (None, ">", "m<"):
        W, V = rc.eigh(H, ["a", "b", "c"], False, sort)
        W0, V0 = rc.eigh(H0, ["a", "b", "c"], False, sort)
        _close(W, W0, 1e-11)
        rec = rc.compose(V, rc.compose(W, V.hc))
        assert rc.almost_equal(to_reference(rec), H0, rtol=1e-10, atol=1e-10)
    assert plugin.hot_calls["qr"] == 1 and plugin.hot_calls["lq"] == 1 and plugin.hot_calls["eigh"] == 3
    assert not plugin.fallbacks


@pytest.mark.parametrize("sym_name,dtype", CASES)
def test_vector_operations_through_cyten_api(plugin, sym_name, dtype):
    A0, B0, A, B = _pair(sym_name, dtype, plugin)
    assert abs(complex(rc.inner(A, B).to_numpy()) - complex(rc.inner(A0, B0).to_numpy())) < 1e-12
    assert abs(float(rc.norm(A).to_numpy()) - float(rc.norm(A0).to_numpy())) < 1e-12
    _close(A + 0.5 * B, A0 + 0.5 * B0, 1e-14)
    _close(A.hc, A0.hc, 0.0)
    for name in ("inner", "norm", "linear_combination", "dagger"):
        assert plugin.hot_calls[name] >= 1, name
    assert not plugin.fallbacks


def test_heff_matvec_matches_reference(plugin):
    """The DMRG two-site effective-Hamiltonian matvec (toycodes/tenpy_toycodes/d_dmrg.py:70-88; 5 permute_legs + 4
    compose) on the synthetic U(1) legs of BASELINE config 1/2: the SAME cyten calls on the reference backend and on
    the plugin."""
    import bench
    chi = 128 if plugin.fast._library is not None else 1024
    H0, th0, ct = bench.build_reference_heff(chi)
    assert ct is rc
    y0 = H0.matvec(th0)
    H, th = bench.build_reference_heff(chi, plugin=plugin)[:2]
    y = H.matvec(th)
    assert y.backend is plugin
    _close(y, y0, 1e-12)
    assert plugin.hot_calls["compose"] == 4 and plugin.hot_calls["permute_legs"] >= 4 and not plugin.fallbacks


def _run_reference_subset(env_extra, report: Path, timeout: int):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([str(REF), str(ROOT / "integration"), str(ROOT)])
    env["CYTEN_B200_REPORT"] = str(report)
    env.update(env_extra)
    cmd = [sys.executable, "-m", "pytest", "tests/python_tests/test_tensors.py", "-p", "pytest_cyten_b200",
           "--block-backends", "torch", "-q", "-p", "no:cacheprovider", "-k", REF_SUBSET]
    r = subprocess.run(cmd, cwd=str(REF), env=env, capture_output=True, text=True, timeout=timeout)
    tail = "\n".join(r.stdout.splitlines()[-25:])
    assert r.returncode == 0, tail + "\n" + r.stderr[-2000:]
    rep = json.loads(report.read_text())
    assert " passed" in tail and " failed" not in tail, tail
    return rep, tail


def _check_report(rep):
    hot = rep["hot_calls"]
    for name in ("compose", "permute_legs", "combine_legs", "split_legs", "svd", "qr", "lq", "eigh",
                 "truncate_singular_values"):
        assert hot.get(name, 0) > 0, (name, rep)
    # the ONLY hot virtual that may fall through is the projection of a leg in the middle of a ChargedTensor's
    # invariant part (hot = first / last leg, as truncated_svd produces them)
    assert set(rep["fallbacks"]) <= {"mask_contract_large_leg"}, rep["fallbacks"]
    assert rep["kernel_launching_calls"] > 1000


@pytest.mark.skipif(not (REF / "tests" / "python_tests" / "test_tensors.py").exists(),
                    reason="reference tests not staged (python oracle/build_ref.py --stage-only)")
def test_reference_pytest_subset_through_the_plugin_oracle(tmp_path):
    rep, tail = _run_reference_subset({"CYTEN_B200_ORACLE": "1"}, tmp_path / "report.json", 900)
    assert rep["library"].startswith("oracle")
    _check_report(rep)


@pytest.mark.gpu
@pytest.mark.skipif(not (REF / "tests" / "python_tests" / "test_tensors.py").exists(),
                    reason="reference tests not staged (python oracle/build_ref.py --stage-only)")
def test_reference_pytest_subset_through_the_plugin_cuda(tmp_path):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    rep, tail = _run_reference_subset({"CYTEN_B200_DEVICE": "cuda:0"}, tmp_path / "report.json", 1800)
    assert rep["library"] == "libcyten_b200.so" and rep["device"].startswith("cuda")
    _check_report(rep)
    out = ROOT / "gpurun_out"
    if out.exists():
        (out / "plugin_reference_subset.json").write_text(json.dumps(dict(report=rep, tail=tail.splitlines()[-3:]), indent=1))

#!/usr/bin/env python3
"""Generates the golden fixtures ``tests/golden/*.npz`` by running the UNMODIFIED reference
(tenpy/cyten built by ``oracle/build_ref.py`` into ``oracle/_ref``) with
``AbelianBackend + NumpyBlockBackend`` on seeded random inputs.

Run in the build container only (needs ``oracle/_ref``):  ``python tests/golden/make_golden.py``.
The fixtures are small (blocks of a few dozen entries) and committed; the GPU box never needs the
reference.  Every fixture stores inputs AND the reference's outputs: block_inds, block contents,
singular values, kept dimensions, truncation error.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT / "oracle" / "_ref"))
import cyten as ct  # noqa: E402
from cyten.testing import randomly_drop_blocks  # noqa: E402

OUT = Path(__file__).resolve().parent


# ---- serialisation ---------------------------------------------------------------------------------
def leg_spec(leg, pre, out):
    if isinstance(leg, ct.AbelianLegPipe):
        out[pre + "kind"] = np.asarray(1)
        out[pre + "is_dual"] = np.asarray(int(leg.is_dual))
        out[pre + "combine_cstyle"] = np.asarray(int(leg.combine_cstyle))
        out[pre + "num_legs"] = np.asarray(len(leg.legs))
        for i, sub in enumerate(leg.legs):
            leg_spec(sub, f"{pre}leg{i}/", out)
        # the reference's tables, to pin the host mirror
        out[pre + "block_ind_map"] = np.asarray(leg.block_ind_map.to_numpy())
        out[pre + "fusion_outcomes_sort"] = np.asarray(leg.fusion_outcomes_sort)
        out[pre + "block_ind_map_slices"] = np.asarray(leg.block_ind_map_slices)
        out[pre + "sector_strides"] = np.asarray(leg.sector_strides)
    else:
        out[pre + "kind"] = np.asarray(0)
        out[pre + "is_dual"] = np.asarray(int(leg.is_dual))
    out[pre + "defining_sectors"] = np.asarray(leg.defining_sectors).reshape(leg.num_sectors, -1)
    out[pre + "multiplicities"] = np.asarray(leg.multiplicities)


def dump_tensor(T, pre, out):
    out[pre + "ncod"] = np.asarray(T.num_codomain_legs)
    out[pre + "ndom"] = np.asarray(T.num_domain_legs)
    for i, l in enumerate(T.codomain):
        leg_spec(l, f"{pre}cod{i}/", out)
    for i, l in enumerate(T.domain):
        leg_spec(l, f"{pre}dom{i}/", out)
    bi = np.asarray(T.data.block_inds.to_numpy()).reshape(len(T.data.blocks), -1)
    out[pre + "block_inds"] = bi
    out[pre + "nblocks"] = np.asarray(len(T.data.blocks))
    for n, b in enumerate(T.data.blocks):
        out[f"{pre}block{n}"] = np.asarray(b.to_numpy())


def dump_diag(D, pre, out):
    leg_spec(D.leg, pre + "leg/", out)
    bi = np.asarray(D.data.block_inds.to_numpy()).reshape(len(D.data.blocks), -1)
    out[pre + "block_inds"] = bi
    out[pre + "nblocks"] = np.asarray(len(D.data.blocks))
    for n, b in enumerate(D.data.blocks):
        out[f"{pre}block{n}"] = np.asarray(b.to_numpy())


# ---- random inputs ---------------------------------------------------------------------------------
SYMS = {
    "u1": (lambda: ct.u1_symmetry, (0,)),
    "z4": (lambda: ct.z4_symmetry, (4,)),
    "u1z3": (lambda: ct.u1_symmetry * ct.z3_symmetry, (0, 3)),
    "u1u1": (lambda: ct.u1_symmetry * ct.u1_symmetry, (0, 0)),
}


def rand_leg(sym, moduli, rng, max_sec=4, max_mult=4, is_dual=None):
    n = int(rng.integers(1, max_sec + 1))
    secs = rng.integers(-2, 3, size=(n, len(moduli)))
    for k, m in enumerate(moduli):
        if m:
            secs[:, k] %= m
    mult = rng.integers(1, max_mult + 1, size=n)
    if is_dual is None:
        is_dual = bool(rng.integers(0, 2))
    return ct.ElementarySpace.from_defining_sectors(sym, secs, mult, is_dual=is_dual)


def rand_tensor(cod, dom, dtype, rng, drop=True, labels=None):
    T = ct.SymmetricTensor.from_random_normal(cod, dom, labels=labels, dtype=dtype)
    if drop and len(T.data.blocks) > 1:
        randomly_drop_blocks(T, max_blocks=None, empty_ok=False, np_random=rng)
    return T


def make_case(symname, dtype_name, seed):
    symf, moduli = SYMS[symname]
    sym = symf()
    rng = np.random.default_rng(seed)
    np.random.seed(seed)
    dtype = ct.Dtype.float64 if dtype_name == "float64" else ct.Dtype.complex128
    out = {"moduli": np.asarray(moduli), "dtype": np.asarray(dtype_name), "seed": np.asarray(seed)}
    for attempt in range(200):
        a, b, c, d, e = (rand_leg(sym, moduli, rng) for _ in range(5))
        A = rand_tensor([a, b], [c, d], dtype, rng, labels=["a", "b", "d*", "c*"])
        B = rand_tensor([c, d], [e], dtype, rng, labels=["c", "d", "e*"])
        if len(A.data.blocks) >= 2 and len(B.data.blocks) >= 1:
            AB = ct.compose(A, B)
            if len(AB.data.blocks) >= 1:
                break
    else:
        raise RuntimeError("no suitable random case")
    dump_tensor(A, "A/", out)
    dump_tensor(B, "B/", out)
    # permute_legs: a deterministic non-trivial permutation mixing codomain and domain
    perm_cod, perm_dom = [2, 0], [3, 1]
    dump_tensor(ct.permute_legs(A, perm_cod, perm_dom), "permute/", out)
    out["permute/codomain"] = np.asarray(perm_cod)
    out["permute/domain"] = np.asarray(perm_dom)
    # compose over two legs, tdot over one leg
    dump_tensor(AB, "compose/", out)
    T1 = ct.tdot(A, B, ["c*"], ["c"])
    dump_tensor(T1, "tdot/", out)
    # combine / split
    C = ct.combine_legs(A, [0, 1], [2, 3])
    dump_tensor(C, "combine/", out)
    dump_tensor(ct.split_legs(C), "split/", out)
    C2 = ct.combine_legs(A, [1, 2] if False else [0, 1])
    dump_tensor(C2, "combine_cod/", out)
    # decompositions
    U, S, Vh = ct.svd(A)
    dump_tensor(U, "svd/U/", out)
    dump_diag(S, "svd/S/", out)
    dump_tensor(Vh, "svd/Vh/", out)
    n_sv = int(sum(len(np.asarray(b.to_numpy())) for b in S.data.blocks))
    chi = max(1, n_sv // 2)
    U2, S2, Vh2, err, renorm = ct.truncated_svd(A, chi_max=chi, svd_min=1e-12)
    dump_tensor(U2, "tsvd/U/", out)
    dump_diag(S2, "tsvd/S/", out)
    dump_tensor(Vh2, "tsvd/Vh/", out)
    out["tsvd/chi_max"] = np.asarray(chi)
    out["tsvd/svd_min"] = np.asarray(1e-12)
    out["tsvd/err"] = np.asarray(err)
    out["tsvd/renormalize"] = np.asarray(renorm)
    Q, R = ct.qr(A)
    dump_tensor(Q, "qr/Q/", out)
    dump_tensor(R, "qr/R/", out)
    H = ct.compose(A, ct.dagger(A))
    dump_tensor(H, "H/", out)
    W, V = ct.eigh(H, new_labels=["x", "y", "z"], new_leg_dual=False)
    dump_diag(W, "eigh/W/", out)
    dump_tensor(V, "eigh/V/", out)
    return out


def appendix_a():
    """The bookkeeping example of SURVEY.md appendix A."""
    np.random.seed(0)
    sym = ct.u1_symmetry
    v = ct.ElementarySpace.from_defining_sectors(sym, [[-1], [0], [1]], [2, 3, 2])
    p = ct.ElementarySpace.from_defining_sectors(sym, [[-1], [1]], [1, 1])
    A = ct.SymmetricTensor.from_random_normal([v, p], [v], labels=["vL", "p", "vR"], dtype=ct.Dtype.float64)
    B = ct.SymmetricTensor.from_random_normal([v, p], [v], labels=["vL", "p", "vR"], dtype=ct.Dtype.float64)
    out = {"moduli": np.asarray((0,)), "dtype": np.asarray("float64"), "seed": np.asarray(0)}
    dump_tensor(A, "A/", out)
    dump_tensor(B, "B/", out)
    th = ct.tdot(A, B, ["vR"], ["vL"])
    dump_tensor(th, "theta/", out)
    dump_tensor(ct.combine_legs(th, [0, 1], [2, 3]), "combine/", out)
    U, S, Vh = ct.svd(th)
    dump_tensor(U, "svd/U/", out)
    dump_diag(S, "svd/S/", out)
    dump_tensor(Vh, "svd/Vh/", out)
    U2, S2, Vh2, err, renorm = ct.truncated_svd(th, chi_max=5, svd_min=1e-12)
    dump_tensor(U2, "tsvd/U/", out)
    dump_diag(S2, "tsvd/S/", out)
    dump_tensor(Vh2, "tsvd/Vh/", out)
    out["tsvd/chi_max"] = np.asarray(5)
    out["tsvd/svd_min"] = np.asarray(1e-12)
    out["tsvd/err"] = np.asarray(err)
    out["tsvd/renormalize"] = np.asarray(renorm)
    return out


def truncation_vectors():
    """Known-answer vectors for _truncate_singular_values_selection (tensor_backend.cpp:138-242)."""
    be = ct.get_backend("abelian", "numpy")
    rng = np.random.default_rng(7)
    out = {}
    n = 0
    for size in (1, 5, 40, 300):
        for kw in (dict(), dict(chi_max=3), dict(chi_max=17, svd_min=1e-3), dict(chi_min=4, svd_min=0.5),
                   dict(degeneracy_tol=0.05, chi_max=11), dict(trunc_cut=0.3), dict(trunc_cut=0.05, chi_max=25),
                   dict(svd_min=1e-8)):
            S = np.sort(rng.random(size) * np.exp(-rng.random(size) * 6))[::-1].copy()
            if size > 10:
                S[size // 2:size // 2 + 3] = S[size // 2]   # an exact degeneracy
                rng.shuffle(S)
            S = S / np.linalg.norm(S)
            args = dict(chi_max=None, chi_min=1, degeneracy_tol=0.0, trunc_cut=0.0, svd_min=None, minimize_error=True)
            args.update(kw)
            mask, err, new_norm = be._truncate_singular_values_selection(
                S, None, args["chi_max"], args["chi_min"], args["degeneracy_tol"], args["trunc_cut"],
                args["svd_min"], args["minimize_error"])
            pre = f"t{n}/"
            out[pre + "S"] = S
            out[pre + "mask"] = np.asarray(mask)
            out[pre + "err"] = np.asarray(err)
            out[pre + "new_norm"] = np.asarray(new_norm)
            out[pre + "args"] = np.asarray([-1 if args["chi_max"] is None else args["chi_max"], args["chi_min"],
                                            args["degeneracy_tol"], args["trunc_cut"],
                                            -1.0 if args["svd_min"] is None else args["svd_min"],
                                            float(args["minimize_error"])], dtype=np.float64)
            n += 1
    out["count"] = np.asarray(n)
    return out


def main():
    np.savez_compressed(OUT / "appendix_a.npz", **appendix_a())
    np.savez_compressed(OUT / "truncation.npz", **truncation_vectors())
    k = 0
    for symname in SYMS:
        for dtype_name in ("float64", "complex128"):
            for seed in (11, 12):
                out = make_case(symname, dtype_name, 1000 * k + seed)
                np.savez_compressed(OUT / f"case_{symname}_{dtype_name}_{seed}.npz", **out)
                k += 1
    print("wrote", len(list(OUT.glob("*.npz"))), "fixtures to", OUT)


if __name__ == "__main__":
    main()

"""numpy model of the two-pass block Jacobi on a rank-deficient matrix: sweeps and work under two rules for pass 1.
  rule "both":   a pair is left alone when BOTH rows are below the noise floor (what factorize.cu does)
  rule "either": a pair is left alone when EITHER row is below the floor; pass 2 then also rotates (significant, noise) pairs
Work is counted in items (block pairs) that rotate (full cost) and items that only form their Gram matrix (0.2 of the cost).
Usage: python tools/rank_deficient_model.py 256 16 128   (n, block, rank)"""
import sys
import numpy as np

EPS = 2.2e-16


def cyclic_pass(G, skip):
    """one cyclic pass of two-sided rotations on G; skip(p, q, a, b) says whether a pair is left alone; returns Q, rotated?"""
    m = G.shape[0]
    Q = np.eye(m)
    G = G.copy()
    rot = False
    for p in range(m - 1):
        for q in range(p + 1, m):
            a, b, g = G[p, p], G[q, q], G[p, q]
            if a <= 0 or b <= 0 or skip(p, q, a, b, g):
                continue
            rot = True
            z = (b - a) / (2 * g)
            t = np.sign(z) / (abs(z) + np.sqrt(1 + z * z)) if z != 0 else 1.0
            c = 1 / np.sqrt(1 + t * t)
            s = c * t
            R = np.array([[c, -s], [s, c]])
            G[[p, q], :] = R @ G[[p, q], :]
            G[:, [p, q]] = G[:, [p, q]] @ R.T
            Q[[p, q], :] = R @ Q[[p, q], :]
    return Q, rot


def sweeps(W, B, rule, pass2, floor2, first_noise_block=0, maxsw=60):
    """runs sweeps until no item rotates; returns (sweeps, work in item-equivalents, W)"""
    n, c = W.shape
    nb = (n + B - 1) // B
    tol = EPS * np.sqrt(c)
    work = 0.0
    for sw in range(maxsw):
        if not pass2:
            W = W[np.argsort(-np.linalg.norm(W, axis=1), kind="stable")]
        any_rot = False
        idx = list(range(nb + (nb & 1)))
        for r in range(len(idx) - 1):
            for k in range(len(idx) // 2):
                i, j = sorted((idx[k], idx[len(idx) - 1 - k]))
                if j >= nb:
                    continue
                if pass2 and j < first_noise_block:
                    continue                                   # (significant, significant): converged in pass 1
                if pass2 and rule == "both" and i < first_noise_block:
                    continue                                   # today's pass 2: noise rows among themselves only
                rows = np.r_[i * B:min((i + 1) * B, n), j * B:min((j + 1) * B, n)]
                X = W[rows]
                G = X @ X.T
                ni = min((i + 1) * B, n) - i * B

                def skip(p, q, a, b, g):
                    # (every pair of the item, as in tools/block_jacobi_model.py: within-block pairs are rotated by every item
                    #  of their block -- more rotations than the kernels do, same convergence behaviour)
                    if g * g <= tol * tol * a * b:
                        return True
                    if pass2:
                        return False
                    return (a <= floor2 and b <= floor2) if rule == "both" else (a <= floor2 or b <= floor2)
                Q, rot = cyclic_pass(G, skip)
                work += 1.0 if rot else 0.2
                if rot:
                    any_rot = True
                    W[rows] = Q @ X
            idx = [idx[0]] + [idx[-1]] + idx[1:-1]              # round-robin: next round's pairing
        if not any_rot:
            return sw + 1, work, W
    return maxsw, work, W


def run(n, B, rank, rule, seed=0):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((n, rank)) @ rng.standard_normal((rank, n))
    floor2 = (EPS * n) ** 2 * np.sum(A * A)
    s1, w1, W = sweeps(A.copy(), B, rule, False, floor2)
    nrm2 = np.sum(W * W, axis=1)
    W = W[np.argsort(-nrm2, kind="stable")]
    nnoise = int(np.sum(np.sort(nrm2)[::-1] <= floor2))
    first = (n - nnoise) // B
    s2, w2, W = sweeps(W, B, rule, True, floor2, first_noise_block=first)
    S = np.sort(np.linalg.norm(W, axis=1))[::-1]
    Vh = W / np.maximum(np.linalg.norm(W, axis=1, keepdims=True), 1e-300)
    ortho = np.abs(Vh @ Vh.T - np.eye(n)).max()
    ref = np.linalg.svd(A, compute_uv=False)
    return dict(rule=rule, pass1_sweeps=s1, pass1_work=round(w1), pass2_sweeps=s2, pass2_work=round(w2), total_work=round(w1 + w2),
                noise_rows=nnoise, s_err=float(np.abs(S - ref).max() / ref[0]), vh_orthogonality=float(ortho))


if __name__ == "__main__":
    n, B, rank = (int(x) for x in sys.argv[1:4])
    for rule in ("both", "either"):
        print(run(n, B, rank, rule), flush=True)

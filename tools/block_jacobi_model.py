"""numpy model of the block one-sided Jacobi (pairs of B-row blocks, k cyclic inner passes on the 2B x 2B Gram matrix).\nUsage: python tools/block_jacobi_model.py 256 16   (round 2: 11 sweeps with 1 inner pass, 10 with 2)"""
import numpy as np, sys
def inner_passes(G, npass, tol):
    # cyclic two-sided Jacobi passes on symmetric G (m x m); returns Q with Q^T G Q more diagonal
    m = G.shape[0]; Q = np.eye(m); G = G.copy()
    for _ in range(npass):
        rot = False
        for p in range(m-1):
            for q in range(p+1, m):
                a,b,g = G[p,p],G[q,q],G[p,q]
                if g*g <= tol*tol*a*b: continue
                rot = True
                z = (b-a)/(2*g); t = np.sign(z)/(abs(z)+np.sqrt(1+z*z)) if z!=0 else 1.0
                c = 1/np.sqrt(1+t*t); s = c*t
                J = np.eye(m); J[p,p]=c; J[q,q]=c; J[p,q]=s; J[q,p]=-s
                # rows: x_p' = c x_p - s x_q ; x_q' = s x_p + c x_q  -> R = [[c,-s],[s,c]]
                R = np.array([[c,-s],[s,c]])
                G[[p,q],:] = R @ G[[p,q],:]; G[:,[p,q]] = G[:,[p,q]] @ R.T
                Q[[p,q],:] = R @ Q[[p,q],:]
        if not rot: break
    return Q
def inner_full(G):
    w,V = np.linalg.eigh(G); return V.T[::-1]
def block_jacobi(A, B=32, npass=1, full=False, maxsw=40):
    W = A.copy(); n = W.shape[0]; nb = n//B; tol = 2.2e-16*np.sqrt(W.shape[1])
    for sw in range(maxsw):
        nr = np.linalg.norm(W,axis=1); W = W[np.argsort(-nr)]
        maxoff = 0
        idx = list(range(nb))
        for r in range(nb-1):
            for k in range(nb//2):
                i,j = idx[k], idx[nb-1-k]
                rows = np.r_[i*B:(i+1)*B, j*B:(j+1)*B]
                X = W[rows]; G = X@X.T
                d = np.sqrt(np.diag(G)); off = np.abs(G)/np.outer(d,d); np.fill_diagonal(off,0); maxoff=max(maxoff,off.max())
                if off.max() <= tol: continue
                Q = inner_full(G) if full else inner_passes(G, npass, tol)
                W[rows] = Q@X
            idx = [idx[0]]+[idx[-1]]+idx[1:-1]
        if maxoff <= tol: return sw+1
    return maxsw
rng = np.random.default_rng(0); n = int(sys.argv[1]); B = int(sys.argv[2])
A = rng.standard_normal((n,n))
for npass in (1,2):
    print("npass",npass, block_jacobi(A,B,npass), flush=True)
print("full", block_jacobi(A,B,full=True))

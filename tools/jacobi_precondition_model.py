"""numpy model: sweeps of the round-robin one-sided Jacobi on A, on R and R^T of A = QR (plain, column-sorted, pivoted), full rank and rank n/2.\nUsage: python tools/jacobi_precondition_model.py 128   (round 2: 12 / 11 / 12 sweeps full rank; 24 -> 12 on R for rank n/2)"""
import numpy as np, sys
def sweeps_onesided(A, tol=1e-14, maxsw=60, sort=True):
    # one-sided Jacobi on ROWS of W (n rows), round-robin parallel ordering, vectorised
    W = A.copy(); n = W.shape[0]
    assert n % 2 == 0
    for sw in range(maxsw):
        if sort:
            nr = np.linalg.norm(W, axis=1); W = W[np.argsort(-nr)]
        idx = np.arange(n)
        maxoff = 0.0
        for r in range(n-1):
            p = idx[:n//2]; q = idx[n//2:][::-1]
            a = np.einsum('ij,ij->i', W[p], W[p]); b = np.einsum('ij,ij->i', W[q], W[q]); g = np.einsum('ij,ij->i', W[p], W[q])
            off = np.abs(g)/np.sqrt(a*b+1e-300)
            maxoff = max(maxoff, off.max())
            rot = off > tol
            zeta = (b-a)/(2*np.where(g==0,1e-300,g))
            t = np.sign(zeta)/(np.abs(zeta)+np.sqrt(1+zeta**2)); t[zeta==0]=1
            c = 1/np.sqrt(1+t*t); s = c*t
            c = np.where(rot,c,1.0); s=np.where(rot,s,0.0)
            Wp = c[:,None]*W[p] - s[:,None]*W[q]; Wq = s[:,None]*W[p] + c[:,None]*W[q]
            W[p]=Wp; W[q]=Wq
            idx = np.concatenate([[idx[0]],[idx[-1]],idx[1:-1]])
        if maxoff < tol: return sw+1
    return maxsw
rng = np.random.default_rng(0)
n = int(sys.argv[1]) if len(sys.argv)>1 else 256
A = rng.standard_normal((n,n))
print("plain A rows:", sweeps_onesided(A))
Q,R = np.linalg.qr(A)
print("R rows (R = upper):", sweeps_onesided(R))
print("R^T rows:", sweeps_onesided(R.T.copy()))
# qr of sorted columns
nr = np.linalg.norm(A,axis=0); Q,R = np.linalg.qr(A[:,np.argsort(-nr)])
print("sortedcol R rows:", sweeps_onesided(R), " R^T rows:", sweeps_onesided(R.T.copy()))
import scipy.linalg as sl
Q,R,P = sl.qr(A, pivoting=True)
print("QRCP R rows:", sweeps_onesided(R), " R^T rows:", sweeps_onesided(R.T.copy()))
Q2,R2 = np.linalg.qr(R.T)
print("QRCP+QR: R2 rows", sweeps_onesided(R2), "R2^T rows", sweeps_onesided(R2.T.copy()))
# rank-deficient
B = rng.standard_normal((n,n//2))@rng.standard_normal((n//2,n))
print("rankdef plain:", sweeps_onesided(B))
Q,R = np.linalg.qr(B); print("rankdef R rows:", sweeps_onesided(R), "R^T rows", sweeps_onesided(R.T.copy()))

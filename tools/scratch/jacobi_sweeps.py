"""How many one-sided Jacobi sweeps does a chunk's Gram need: W = G (what K3 does) against W = L^T
(pivoted Cholesky factor, G = L L^T), same round-robin ordering, same stop rule (max cosine <= tol)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from oracle.swfd_batched_oracle import BatchedSeqSWFD

def round_robin(n):
    idx = list(range(n))
    for _ in range(n - 1):
        yield [(idx[i], idx[n - 1 - i]) for i in range(n // 2)]
        idx = [idx[0]] + [idx[-1]] + idx[1:-1]

def sweeps_needed(W, tol=1e-3, max_sweeps=30):
    W = W.copy()
    n = W.shape[1]
    hist = []
    for sw in range(max_sweeps):
        off = 0.0
        for pairs in round_robin(n):
            p = np.array([a for a, b in pairs]); q = np.array([b for a, b in pairs])
            x, y = W[:, p], W[:, q]
            g = np.einsum("ij,ij->j", x, y); nx = np.einsum("ij,ij->j", x, x); ny = np.einsum("ij,ij->j", y, y)
            prod = nx * ny
            nm = max(nx.max(), ny.max())
            live = prod > 1e-24 * nm * nm
            rel = np.where(live, np.abs(g) / np.sqrt(np.where(live, prod, 1.0)), 0.0)
            off = max(off, rel.max())
            a, b = ny - nx, 2 * g
            r = np.sqrt(a * a + b * b)
            t = np.where(rel > 1e-10, b / np.where(np.abs(a) + r > 0, np.abs(a) + r, 1.0) * np.where(a < 0, -1.0, 1.0), 0.0)
            c = 1 / np.sqrt(1 + t * t); s = c * t
            W[:, p], W[:, q] = c * x - s * y, s * x + c * y
        hist.append(off)
        if off <= tol:
            break
    return len(hist), hist, W

def pivoted_cholesky(G, rtol=1e-9):
    G = G.copy(); n = G.shape[0]
    L = np.zeros((n, n)); perm = np.arange(n)
    d = np.diag(G).copy(); d0 = d.max()
    for k in range(n):
        j = k + np.argmax(d[perm[k:]])
        perm[[k, j]] = perm[[j, k]]
        pk = perm[k]
        if d[pk] <= rtol * d0:
            break
        L[pk, k] = np.sqrt(d[pk])
        rest = perm[k + 1:]
        L[rest, k] = (G[rest, pk] - L[rest, :k] @ L[pk, :k]) / L[pk, k]
        d[rest] -= L[rest, k] ** 2
    return L

ell, d, N = 128, 1024, 10000
A = synth.lowrank_stream(3000, d, seed=0, scale=2.0).astype(np.float64)
R = synth.row_norm2_max(A.astype(np.float32))
ora = BatchedSeqSWFD(N, R, d, ell)
ora.fit(A[:2560])
X = A[2560:2560 + ell]
for j in (0, 2, 4):
    side = ora.fd[j]
    thr = ora.thresholds[j]
    raw = np.einsum("ij,ij->i", X, X) >= thr
    S = np.vstack([side.sketch, X[~raw]])
    G = S @ S.T
    G += 1e-7 * np.abs(G).max() * np.random.default_rng(1).standard_normal(G.shape)  # 3xTF32-level noise
    G = 0.5 * (G + G.T)
    n1, h1, W1 = sweeps_needed(G)
    lam_ref = np.sort(np.linalg.eigvalsh(G))[::-1]
    lam1 = np.sort(np.linalg.norm(W1, axis=0))[::-1]
    L = pivoted_cholesky(G)
    n2, h2, W2 = sweeps_needed(L)            # columns of L: G = L L^T, left singular vectors = eigenvectors
    lam2 = np.sort(np.einsum("ij,ij->j", W2, W2))[::-1]
    n3, h3, W3 = sweeps_needed(np.ascontiguousarray(L.T))   # rows of L
    lam3 = np.sort(np.einsum("ij,ij->j", W3, W3))[::-1]
    print("level %d rows %d: W=G %d sweeps (err %.1e) | W=L %d sweeps (err %.1e) | W=L^T %d sweeps (err %.1e)" % (
        j, S.shape[0], n1, np.abs(lam1 - lam_ref[:len(lam1)]).max() / lam_ref[0],
        n2, np.abs(lam2 - np.maximum(lam_ref, 0)[:len(lam2)]).max() / lam_ref[0],
        n3, np.abs(lam3 - np.maximum(lam_ref, 0)[:len(lam3)]).max() / lam_ref[0]))
    print("   off history G:", ["%.1e" % v for v in h1])
    print("   off history L:", ["%.1e" % v for v in h2])
    print("   off history L^T:", ["%.1e" % v for v in h3])

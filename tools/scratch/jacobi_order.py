import sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tools/scratch")
from mused_b200 import synth
from oracle.swfd_batched_oracle import BatchedSeqSWFD

def odd_even_pairs(n):
    # odd-even transposition on positions: round r pairs (i, i+1) for i = r%2, r%2+2, ..; columns trade places
    for r in range(n):
        yield [(i, i + 1) for i in range(r % 2, n - 1, 2)]

def run(G, mode, tol=1e-3):
    W = G.copy(); n = W.shape[1]
    if mode in ("sort", "sort+rijk"):
        W = W[:, np.argsort(-np.linalg.norm(W, axis=0))]
    hist = []
    for sw in range(30):
        off = 0.0
        for pairs in odd_even_pairs(n):
            p = np.array([a for a, b in pairs]); q = np.array([b for a, b in pairs])
            x, y = W[:, p], W[:, q]
            g = np.einsum("ij,ij->j", x, y); nx = np.einsum("ij,ij->j", x, x); ny = np.einsum("ij,ij->j", y, y)
            prod = nx * ny; nm = max(nx.max(), ny.max())
            live = prod > 1e-24 * nm * nm; pos = prod > 0
            rel = np.where(pos, np.abs(g) / np.sqrt(np.where(pos, prod, 1.0)), 0.0)
            off = max(off, rel[live].max() if live.any() else 0.0)
            a, b = ny - nx, g + g
            r = np.sqrt(a * a + b * b); den = np.abs(a) + r
            t = np.where(pos & (rel > 1e-10), b / np.where(den > 0, den, 1.0) * np.where(a < 0, -1.0, 1.0), 0.0)
            c = 1 / np.sqrt(1 + t * t); s = c * t
            xn, yn = c * x - s * y, s * x + c * y
            # odd-even transposition: the two columns trade places after meeting
            if mode in ("rijk", "sort+rijk"):
                nxn = np.einsum("ij,ij->j", xn, xn); nyn = np.einsum("ij,ij->j", yn, yn)
                keep = nxn >= nyn          # larger norm moves to the LEFT position always (sorting network)
                W[:, p] = np.where(keep, xn, yn); W[:, q] = np.where(keep, yn, xn)
            else:
                W[:, p], W[:, q] = yn, xn
        hist.append(off)
        if off <= tol: break
    return hist, W

ell, d, N = 128, 1024, 10000
A = synth.lowrank_stream(3000, d, seed=0, scale=2.0).astype(np.float64)
R = synth.row_norm2_max(A.astype(np.float32))
ora = BatchedSeqSWFD(N, R, d, ell)
ora.fit(A[:2560])
X = A[2560:2560 + ell]
for j in (1, 3, 5):
    side = ora.fd[j]
    raw = np.einsum("ij,ij->i", X, X) >= ora.thresholds[j]
    S = np.vstack([side.sketch, X[~raw]])
    if S.shape[0] % 2: S = np.vstack([S, np.zeros((1, d))])
    G = S @ S.T
    G += 1e-7 * np.abs(G).max() * np.random.default_rng(1).standard_normal(G.shape)
    G = 0.5 * (G + G.T)
    lam_ref = np.sort(np.linalg.eigvalsh(G))[::-1]
    for mode in ("plain", "sort", "rijk", "sort+rijk"):
        h, W = run(G, mode)
        lam = np.sort(np.linalg.norm(W, axis=0))[::-1]
        print("level %d %-9s %2d sweeps  %s  err %.1e" % (j, mode, len(h), " ".join("%.0e" % v for v in h), np.abs(lam - np.abs(lam_ref)).max() / lam_ref[0]), flush=True)

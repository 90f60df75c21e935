"""Emulation of the mixed-precision K3: fp32 one-sided Jacobi on W = G / gmax until the largest cosine is
<= tol32, V32 = normalised columns, W = G V32 in float64, float64 sweeps with the kernel's stop rule."""
import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from oracle.swfd_batched_oracle import BatchedSeqSWFD
sys.path.insert(0, "tools/scratch")
from jacobi_sweeps import round_robin

def sweep(W, dt, tol_apply):
    n = W.shape[1]
    off = dt(0)
    for pairs in round_robin(n):
        p = np.array([a for a, b in pairs]); q = np.array([b for a, b in pairs])
        x, y = W[:, p], W[:, q]
        g = np.einsum("ij,ij->j", x, y, dtype=dt); nx = np.einsum("ij,ij->j", x, x, dtype=dt); ny = np.einsum("ij,ij->j", y, y, dtype=dt)
        prod = nx * ny
        nm = max(nx.max(), ny.max())
        live = prod > dt(1e-24) * nm * nm if dt == np.float64 else prod > dt(1e-20) * nm * nm
        pos = prod > 0
        rel = np.where(pos, np.abs(g) / np.sqrt(np.where(pos, prod, dt(1))), dt(0)).astype(dt)
        off = max(off, rel[live].max() if live.any() else dt(0))
        a, b = ny - nx, g + g
        r = np.sqrt(a * a + b * b)
        den = np.abs(a) + r
        t = np.where(pos & (rel > tol_apply), b / np.where(den > 0, den, dt(1)) * np.where(a < 0, dt(-1), dt(1)), dt(0)).astype(dt)
        c = (dt(1) / np.sqrt(dt(1) + t * t)).astype(dt); s = (c * t).astype(dt)
        W[:, p], W[:, q] = (c * x - s * y).astype(dt), (s * x + c * y).astype(dt)
    return float(off)

if __name__ == "__main__":
    ell, d, N = 128, 1024, 10000
    A = synth.lowrank_stream(3000, d, seed=0, scale=2.0).astype(np.float64)
    R = synth.row_norm2_max(A.astype(np.float32))
    ora = BatchedSeqSWFD(N, R, d, ell)
    ora.fit(A[:2560])
    X = A[2560:2560 + ell]
    for j in (1, 2, 3, 4, 5):
        side = ora.fd[j]
        raw = np.einsum("ij,ij->i", X, X) >= ora.thresholds[j]
        S = np.vstack([side.sketch, X[~raw]])
        if S.shape[0] % 2: S = np.vstack([S, np.zeros((1, d))])
        G = S @ S.T
        G += 1e-7 * np.abs(G).max() * np.random.default_rng(1).standard_normal(G.shape)
        G = 0.5 * (G + G.T)
        lam_ref = np.sort(np.linalg.eigvalsh(G))[::-1]
        # float64 only
        W = G.copy(); h64 = []
        for sw in range(30):
            o = sweep(W, np.float64, 1e-10); h64.append(o)
            if o <= 1e-3: break
        # hybrid
        W32 = (G / np.abs(np.diag(G)).max()).astype(np.float32); h32 = []
        for sw in range(20):
            o = sweep(W32, np.float32, np.float32(1e-7)); h32.append(o)
            if o <= 1e-3: break
        V = W32.astype(np.float64)
        nrm = np.linalg.norm(V, axis=0)
        V = np.where(nrm > 0, V / np.where(nrm > 0, nrm, 1), 0.0)
        orth = np.abs(V.T @ V - np.eye(V.shape[1]))[np.ix_(nrm > 1e-6 * nrm.max(), nrm > 1e-6 * nrm.max())].max()
        W = G @ V; hh = []
        for sw in range(10):
            o = sweep(W, np.float64, 1e-10); hh.append(o)
            if o <= 1e-3: break
        lam = np.sort(np.linalg.norm(W, axis=0))[::-1]
        big = lam_ref > 1e-3 * lam_ref[0]
        print("level %d n=%d: f64 %d sweeps | hybrid %d fp32 %s + %d f64 %s ; |V^T V - I| %.1e ; eig rel err (lam > 1e-3 max) %.1e, abs/max %.1e" % (
            j, G.shape[0], len(h64), len(h32), ["%.0e" % v for v in h32[-3:]], len(hh), ["%.0e" % v for v in hh], orth,
            np.max(np.abs(lam[big] - lam_ref[big]) / lam_ref[big]), np.max(np.abs(lam - np.abs(lam_ref))) / lam_ref[0]), flush=True)

"""NumPy prototype of the DEVICE algorithm (not the oracle, not product code).

The CUDA engine (mused_b200/csrc/swfd_engine.cu) replaces the reference's
"thin SVD of [sketch; buffer]" (swfd/frequent_directions.py:39-58) by an
incrementally maintained eigen-decomposition of the Gram matrix of the stored
rows: every incoming row borders the (diagonalised) Gram with one row/column,
which is a symmetric ARROWHEAD eigenproblem solved through its secular
equation (Gu & Eisenstat style: roots kept as origin+offset, coupling vector
recomputed from the roots so eigenvectors are orthogonal to working precision).
This file states that algorithm in NumPy, step for step as the kernel does it,
so that its numerics can be checked on the CPU against the reference before a
GPU minute is spent.  tests/test_proto_arrowhead.py runs it.
"""
from __future__ import annotations

import numpy as np

EPS = np.finfo(np.float64).eps


def arrowhead_eig(d, w, alpha, max_iter=80):
    """Eigen-decomposition of M = [[diag(d), w], [w^T, alpha]].

    Returns (lam, U) with M = U diag(lam) U^T, lam NOT sorted: the first
    len(d) entries correspond to ... see ``order`` below; callers sort.
    d may be unsorted and may contain equal entries; w may contain zeros.
    """
    d = np.asarray(d, dtype=np.float64).copy()
    w = np.asarray(w, dtype=np.float64).copy()
    n = d.shape[0]
    tip = n
    G = np.eye(n + 1)                       # accumulated type-2 rotations
    scale = max(np.max(np.abs(d), initial=0.0), abs(alpha), np.max(np.abs(w), initial=0.0))
    tol = 8.0 * EPS * scale
    nd = np.abs(w) > tol                    # type-1 deflation
    w[~nd] = 0.0
    # ---- type-2 deflation: (nearly) equal diagonal entries ------------------
    idx = np.where(nd)[0]
    idx = idx[np.argsort(-d[idx], kind="stable")]
    if len(idx) > 1:
        cur = idx[0]
        for nxt in idx[1:]:
            r = np.hypot(w[cur], w[nxt])
            c = w[nxt] / r
            s = w[cur] / r
            t = d[cur] - d[nxt]
            if abs(t * c * s) <= tol:
                # rotate plane (cur, nxt) so that w[cur] -> 0, w[nxt] -> r
                # new basis: e_cur' = c e_cur - s e_nxt ; e_nxt' = s e_cur + c e_nxt
                gc, gn = G[:, cur].copy(), G[:, nxt].copy()
                G[:, cur] = c * gc - s * gn
                G[:, nxt] = s * gc + c * gn
                dc, dn = d[cur], d[nxt]
                d[cur] = dc * c * c + dn * s * s
                d[nxt] = dc * s * s + dn * c * c
                w[cur] = 0.0
                w[nxt] = r
                nd[cur] = False
            cur = nxt
    idx = np.where(nd)[0]
    idx = idx[np.argsort(-d[idx], kind="stable")]
    m = len(idx)
    lam = np.empty(n + 1)
    U = np.zeros((n + 1, n + 1))
    # deflated eigenpairs pass through
    for i in np.where(~nd)[0]:
        lam[i] = d[i]
        U[i, i] = 1.0
    if m == 0:
        lam[tip] = alpha
        U[tip, tip] = 1.0
        return lam, G @ U
    dd = d[idx]
    ww = w[idx]
    w2 = ww * ww
    wn = np.sqrt(np.sum(w2))
    # ---- secular roots: m+1 of them, root k kept as org[k] + mu[k] ----------
    org = np.empty(m + 1, dtype=np.int64)   # index into dd of the origin pole
    mu = np.empty(m + 1)
    for k in range(m + 1):
        if k == 0:
            o = 0
            a, b = 0.0, max(dd[0], alpha) + wn - dd[0]
            a_pole, b_pole = True, False
        elif k == m:
            o = m - 1
            a, b = min(dd[m - 1], alpha) - wn - dd[m - 1], 0.0
            a_pole, b_pole = False, True
        else:
            gap = dd[k - 1] - dd[k]
            half = 0.5 * gap
            delta = dd - dd[k]
            hm = (alpha - dd[k]) - half - np.sum(w2 / (delta - half))
            if hm > 0:                      # root in the upper half: origin = upper pole
                o = k - 1
                a, b = -half, 0.0
                a_pole, b_pole = False, True
            else:
                o = k
                a, b = 0.0, half
                a_pole, b_pole = True, False
            if hm == 0:
                org[k] = k
                mu[k] = half
                continue
        delta = dd - dd[o]
        c0 = alpha - dd[o]

        def h_parts(x):
            den = delta - x
            q = w2 / den
            return c0 - x - np.sum(q), -1.0 - np.sum(q / den), abs(c0) + abs(x) + np.sum(np.abs(q))

        # pole-removed function F = p(x) * h(x), p vanishing at the pole end(s)
        # of the bracket; F is smooth on the closed bracket.
        x = 0.5 * (a + b)
        if not a_pole and not b_pole:
            pass
        for it in range(max_iter):
            h, hp, hs = h_parts(x)
            if h == 0 or abs(h) <= 8.0 * EPS * hs:
                break
            if h > 0:
                a = x
                a_pole = False
            else:
                b = x
                b_pole = False
            # Newton on F(x) = (x - pa)(pb - x) h(x) with the poles that are
            # still bracket ends; falls back to bisection when it leaves (a,b)
            pa = 0.0 if (o == k and k < m) or (k == 0) else None      # pole at origin below root
            # origin pole sits at x=0; it is the lower end when origin==k (root above pole)
            # and the upper end when origin==k-1 (root below pole)
            lower_pole0 = (k == 0) or (k < m and o == k)
            if lower_pole0:
                p, pp = x, 1.0
            else:
                p, pp = -x, -1.0
            F = p * h
            Fp = pp * h + p * hp
            xn = x - F / Fp if Fp != 0 else 0.5 * (a + b)
            if not (a < xn < b) or not np.isfinite(xn):
                xn = 0.5 * (a + b)
            if xn == x or (b - a) <= 2.0 * EPS * max(abs(a), abs(b)):
                x = xn
                break
            x = xn
        org[k] = o
        mu[k] = x
    # lam_k - dd_i computed without cancellation
    diff = (dd[org][:, None] - dd[None, :]) + mu[:, None]      # (m+1) x m : lam_k - d_i
    # ---- Gu-Eisenstat: coupling vector for which the computed roots are exact
    what = np.empty(m)
    for i in range(m):
        prod = -diff[i, i] * diff[i + 1, i]
        for j in range(m):
            if j < i:
                prod *= diff[j, i] / (dd[j] - dd[i])
            elif j > i:
                prod *= diff[j + 1, i] / (dd[j] - dd[i])
        what[i] = np.sign(ww[i]) * np.sqrt(max(prod, 0.0))
    # ---- eigenvectors --------------------------------------------------------
    for k in range(m + 1):
        col = np.zeros(n + 1)
        top = what / (-diff[k, :])          # what_i / (d_i - lam_k)
        nrm = np.sqrt(1.0 + np.sum(top * top))
        col[idx] = top / nrm
        col[tip] = -1.0 / nrm
        lamk = dd[org[k]] + mu[k]
        # slot: roots 0..m-1 take the slots of the non-deflated d's, root m the tip
        slot = idx[k] if k < m else tip
        lam[slot] = lamk
        U[:, slot] = col
    return lam, G @ U


class ArrowFD:
    """Drop-in for oracle.swfd_oracle.FDDump computing the same quantities the
    way the CUDA engine does: rows R (n x d), Gram model R R^T = Q diag(lam) Q^T."""

    def __init__(self, d, sketch_dim, error):
        assert d > sketch_dim, "engine requires d > sketch_dim"
        self.d, self.ell, self.error = d, sketch_dim, error
        self.R = np.zeros((self.ell, d))
        self.lam = np.zeros(self.ell)
        self.Q = np.eye(self.ell)
        self.delta = 0.0
        self.energy = 0.0
        self.b = 0
        self.s2 = np.zeros(self.ell)
        self.n_flush = 0

    @property
    def pending(self):
        return None if self.b == 0 else self.R[self.ell:]

    def _border(self, x):
        g = self.R @ x
        alpha = float(x @ x)
        w = self.Q.T @ g
        lam, U = arrowhead_eig(self.lam, w, alpha)
        n = self.lam.shape[0]
        Qe = np.zeros((n + 1, n + 1))
        Qe[:n, :n] = self.Q
        Qe[n, n] = 1.0
        Qn = Qe @ U
        order = np.argsort(-lam, kind="stable")
        self.lam = lam[order]
        self.Q = Qn[:, order]
        self.R = np.vstack([self.R, x[None, :]])
        self.b += 1

    def flush(self):
        if self.b == 0:
            return
        ell = self.ell
        lam = np.maximum(self.lam, 0.0)
        cut = lam[ell]
        self.delta += cut / 2.0
        s2 = lam[:ell] - cut
        with np.errstate(divide="ignore", invalid="ignore"):
            scale = np.where(lam[:ell] > 0, np.sqrt(s2 / np.where(lam[:ell] > 0, lam[:ell], 1.0)), 0.0)
        self.R = (self.Q[:, :ell] * scale[None, :]).T @ self.R
        self.lam = s2.copy()
        self.s2 = s2.copy()
        self.Q = np.eye(ell)
        self.b = 0
        self.energy = self.s2[0]
        self.n_flush += 1

    def fit(self, X):
        x = np.asarray(X, dtype=np.float64).reshape(-1)
        self.energy = self.energy + float(x @ x)
        self._border(x)
        if self.b >= self.ell:
            self.flush()
        elif self.energy >= self.error:
            self.flush()

    def dump(self):
        if self.b == 0 and self.s2[0] >= self.error:
            v = self.R[0:1].copy()
            self.R[0, :] = 0
            self.s2[0] = 0
            self.lam[0] = 0
            self.energy = 0.0
            return v
        if self.b != 0 and self.s2[0] >= self.error:
            raise AssertionError("dump with pending rows and s2[0] >= error cannot happen")
        return None

    @property
    def sketch(self):
        return self.R[:self.ell]

    def get_sketch(self):
        self.flush()
        return self.R

    def get(self):
        self.flush()
        with np.errstate(divide="ignore", invalid="ignore"):
            vt = np.where(self.s2[:, None] > 0, self.R / np.sqrt(np.where(self.s2 > 0, self.s2, 1.0))[:, None], 0.0)
        return self.R, self.s2, vt, self.delta

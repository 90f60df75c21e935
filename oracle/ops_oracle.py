"""NumPy (float64) restatements of the similarity / assignment operators.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

``cosine_assign`` is the north_star operator; the reference has no such
function, so its parity is UNPINNED by any reference
test -- the restatement follows sklearn's cosine_similarity semantics used at
matrix_operations.py:106 (rows L2-normalised, zero rows keep norm 1) plus
np.argmax, and the nearest-centroid rule of matrix_operations.py:286-289.
``knn_adjacency`` follows matrix_operations.py:22-53 / :101-108 / :112-119 / :123-130 and
is checked against the imported reference by tests/test_oracle_golden.py.
"""
from __future__ import annotations

import numpy as np


def _normalize(X):
    n = np.linalg.norm(X, axis=1, keepdims=True)
    n[n == 0] = 1.0
    return X / n


def cosine_assign(X, C, tau):
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    S = _normalize(X) @ _normalize(C).T
    j = np.argmax(S, axis=1)
    best = S[np.arange(len(X)), j]
    labels = np.where(best >= tau, j, -1).astype(np.int32)
    srt = np.sort(S, axis=1)
    gap = srt[:, -1] - srt[:, -2] if S.shape[1] > 1 else np.full(len(X), np.inf)
    return labels, best, gap


def l2_assign(X, C, tau):
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    D = np.sqrt(np.maximum((X * X).sum(1)[:, None] + (C * C).sum(1)[None, :] - 2 * X @ C.T, 0))
    j = np.argmin(D, axis=1)
    best = D[np.arange(len(X)), j]
    labels = np.where((tau < 0) | (best <= tau), j, -1).astype(np.int32)
    srt = np.sort(D, axis=1)
    gap = srt[:, 1] - srt[:, 0] if D.shape[1] > 1 else np.full(len(X), np.inf)
    return labels, best, gap


def haversine_km(F):
    """Pairwise haversine distance of (latitude, longitude) rows in degrees,
    matrix_operations.py:250-263 (r = 6371 km)."""
    lat = np.radians(F[:, 0])
    lon = np.radians(F[:, 1])
    dlat = lat[None, :] - lat[:, None]
    dlon = lon[None, :] - lon[:, None]
    a = np.sin(dlat / 2) ** 2 + np.cos(lat)[:, None] * np.cos(lat)[None, :] * np.sin(dlon / 2) ** 2
    return 2 * np.arcsin(np.sqrt(np.clip(a, 0.0, 1.0))) * 6371


def knn_keys(F, metric):
    F = np.asarray(F, dtype=np.float64)
    if metric == "cosine":
        Fn = _normalize(F)
        return Fn @ Fn.T
    if metric == "haversine":                                  # matrix_operations.py:22-30
        return -haversine_km(F)
    if metric == "l1sum":                                      # matrix_operations.py:40-50
        K = np.zeros((len(F), len(F)))
        for j in range(F.shape[1]):
            K -= np.abs(F[None, :, j] - F[:, None, j])
        return K
    sq = (F * F).sum(1)
    K = -np.maximum(sq[:, None] + sq[None, :] - 2 * F @ F.T, 0)
    np.fill_diagonal(K, 0.0)
    return K


def knn_adjacency(F, k, metric):
    """Directed 0/1 adjacency; returns (adj, margin) where margin[i] is the key gap
    between the k-th and (k+1)-th neighbour (near-tie detector)."""
    K = knn_keys(F, metric)
    W = K.shape[0]
    k = min(k, W)
    order = np.argsort(-K, axis=1, kind="stable")
    adj = np.zeros((W, W), dtype=np.uint8)
    for i in range(W):
        for j in order[i, :k]:
            if j != i:
                adj[i, j] = 1
    srt = -np.sort(-K, axis=1)
    margin = srt[:, k - 1] - srt[:, k] if k < W else np.full(W, np.inf)
    return adj, margin


def contingency(prev, cur, P, Q):
    out = np.zeros((P, Q), dtype=np.int32)
    for p, q in zip(prev, cur):
        if 0 <= p < P and 0 <= q < Q:
            out[p, q] += 1
    return out


def username_adjacency(col):
    """matrix_operations.py:55-72 + writer :123-130: rows with the same non-empty user name are neighbours."""
    W = len(col)
    out = np.zeros((W, W))
    valid = [i for i in range(W) if col[i] != ""]
    groups = {}
    for r, i in enumerate(valid):
        groups.setdefault(col[i], []).append(r)
    for r, i in enumerate(valid):
        for q in groups[col[i]]:
            if valid[q] != i:
                out[i, valid[q]] = 1
    return out


def jaccard_matrix(col):
    """:74-89 similarities of the valid rows (jaccard_similarity :245-248, -1 on the diagonal) and their indices."""
    W = len(col)
    valid = [i for i in range(W) if col[i] != ""]
    sets = [set(col[i]) if col[i] else set() for i in valid]
    n = len(valid)
    S = np.zeros((n, n))
    for a in range(n):
        for b in range(n):
            if a == b:
                S[a, b] = -1
            elif sets[a] and sets[b]:
                S[a, b] = len(sets[a] & sets[b]) / len(sets[a] | sets[b])
    return S, np.asarray(valid)


def tags_adjacency(col, k):
    """The tags branch with ties broken by the lowest index (np.argsort(kind="stable")); the reference's default
    argsort leaves the order inside ties to numpy's introsort, so only rows whose k-th and (k+1)-th similarities
    differ are determined by the reference itself.  Returns (adjacency, mask of such determined rows)."""
    S, valid = jaccard_matrix(col)
    W = len(col)
    out = np.zeros((W, W))
    determined = np.ones(W, dtype=bool)
    for a in range(len(valid)):
        order = np.argsort(-S[a], kind="stable")
        top = order[:k]
        if k < len(valid) and S[a, order[k - 1]] == S[a, order[k]]:
            determined[valid[a]] = False
        for b in top:
            if valid[b] != valid[a]:
                out[valid[a], valid[b]] = 1
    return out, determined

"""Numeric operators of /root/reference/matrix_operations.py on the B200.

Same names and argument meaning as the reference for the functions main.py
imports.  The string modalities "username" and "tags" (matrix_operations.py:55-89)
map their strings to integer ids on the host (dictionaries) and do the pairwise
work on the device; the TF-IDF vocabulary of the "text" branch (:91-100) stays
host string processing (sklearn) -- its matrix goes through the device cosine
top-k like any dense modality ("text" / "cosine").
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

_METRIC = {"cosine": _lib.METRIC_COSINE, "text": _lib.METRIC_COSINE,
           "euclidean": _lib.METRIC_L2, "l2": _lib.METRIC_L2}
_METRIC_F64 = {"haversine": _lib.METRIC_HAVERSINE, "location": _lib.METRIC_HAVERSINE,
               "l1sum": _lib.METRIC_L1SUM, "time": _lib.METRIC_L1SUM}


def _dev_f32(x, device):
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32).contiguous()
    return torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(device)


def cosine_assign(X, centroids, threshold, *, metric="cosine", device="cuda:0", return_score=False):
    """Fused similarity + threshold + argmax event assignment (north_star operator).

    label[i] = argmax_j cos(X[i], centroids[j]) if that cosine >= threshold else -1
    (metric="l2": argmin distance, kept when distance <= threshold, threshold < 0 = always).
    Nearest reference code: matrix_operations.py:106 (cosine_similarity) and :286-289 (argmin).
    Returns int32 labels (torch tensor on ``device``) [and float32 scores].
    """
    lib = _lib.load()
    Xd, Cd = _dev_f32(X, device), _dev_f32(centroids, device)
    if Xd.dim() != 2 or Cd.dim() != 2 or Xd.shape[1] != Cd.shape[1]:
        raise ValueError("X (n x d) and centroids (k x d) must share d")
    n, d = Xd.shape
    labels = torch.empty(n, dtype=torch.int32, device=Xd.device)
    # labels only -> NULL score: the library takes its screening + re-scoring path (bandwidth-bound, same labels)
    score = torch.empty(n, dtype=torch.float32, device=Xd.device) if return_score else None
    with torch.cuda.device(Xd.device):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.mused_assign(Xd.data_ptr(), n, d, Xd.stride(0) if n > 1 else d, Cd.data_ptr(),
                                    Cd.shape[0], Cd.stride(0) if Cd.shape[0] > 1 else d,
                                    _METRIC[metric], float(threshold), labels.data_ptr(),
                                    score.data_ptr() if return_score else None, st))
    return (labels, score) if return_score else labels


def topk_adjacency(features, k, *, metric="cosine", out=None, device="cuda:0"):
    """kNN adjacency of one dense modality, OR-accumulated into ``out`` (uint8 W x W).

    cosine: matrix_operations.py:101-108 (k = k_basis + 1 there); euclidean:
    :112-119 (k = k_basis); haversine ("location", :22-30, rows = latitude,
    longitude in degrees) and l1sum ("time", :32-53, rows = datetaken,
    dateupload) are computed in float64; writer :123-130 (directed, no self edges).
    """
    lib = _lib.load()
    if metric in _METRIC_F64:
        # degrees and epoch seconds need float64 (matrix_operations.py:22-53)
        if isinstance(features, torch.Tensor):
            F = features.to(device=device, dtype=torch.float64).contiguous()
        else:
            F = torch.as_tensor(np.ascontiguousarray(features, dtype=np.float64)).to(device)
        W, f = F.shape
        if out is None:
            out = torch.zeros((W, W), dtype=torch.uint8, device=F.device)
        with torch.cuda.device(F.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(lib.mused_topk_adjacency_f64(F.data_ptr(), W, f, F.stride(0) if W > 1 else f,
                                                    _METRIC_F64[metric], int(k), out.data_ptr(), st))
        return out
    F = _dev_f32(features, device)
    W, f = F.shape
    if out is None:
        out = torch.zeros((W, W), dtype=torch.uint8, device=F.device)
    with torch.cuda.device(F.device):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.mused_topk_adjacency(F.data_ptr(), W, f, F.stride(0) if W > 1 else f,
                                            _METRIC[metric], int(k), out.data_ptr(), st))
    return out


STRING_MODALITIES = ("username", "tags")


def _first_column(data):
    a = np.asarray(data, dtype=object)
    return a[:, 0] if a.ndim == 2 else a


def _string_adjacency(data, modality_type, k_basis, out, device):
    """The "username" (matrix_operations.py:55-72) and "tags" (:74-89) branches: strings -> integer ids on the
    host (dictionaries), all pairwise work + top-k + edge writing on the device; edges are OR-ed into ``out``."""
    lib = _lib.load()
    col = _first_column(data)
    W = len(col)
    valid = np.array([i for i in range(W) if not (isinstance(col[i], str) and col[i] == "")], dtype=np.int64)   # data[:, 0] != ''
    if len(valid) == 0:
        return out
    nv = len(valid)
    sub = out if nv == W else torch.zeros((nv, nv), dtype=torch.uint8, device=out.device)
    with torch.cuda.device(out.device):
        st = torch.cuda.current_stream().cuda_stream
        if modality_type == "username":
            ids, table = np.empty(nv, dtype=np.int32), {}
            for r, i in enumerate(valid):
                ids[r] = table.setdefault(col[i], len(table))
            g = torch.as_tensor(ids).to(out.device)
            _lib.check(lib.mused_group_adjacency(g.data_ptr(), nv, sub.data_ptr(), st))
        else:
            # set(tags): the characters of a string, the items of a list (:83)
            table, rows = {}, []
            for i in valid:
                t = col[i]
                rows.append(sorted({table.setdefault(tok, len(table)) for tok in t}) if t is not None and len(t) else [])
            offs = np.zeros(nv + 1, dtype=np.int32)
            offs[1:] = np.cumsum([len(r) for r in rows])
            flat = np.fromiter((v for r in rows for v in r), dtype=np.int32, count=int(offs[-1]))
            o = torch.as_tensor(offs).to(out.device)
            f = torch.as_tensor(flat if len(flat) else np.zeros(1, dtype=np.int32)).to(out.device)
            _lib.check(lib.mused_jaccard_topk_adjacency(o.data_ptr(), f.data_ptr(), nv, min(int(k_basis), nv),
                                                        sub.data_ptr(), st))
    if sub is not out:
        idx = torch.as_tensor(valid, device=out.device)
        out[idx[:, None], idx[None, :]] |= sub
    return out


def create_adjacency_matrix(data, modality_type, k_basis=50, *, device="cuda:0"):
    """create_adjacency_matrix of matrix_operations.py:14-132 -> W x W float64 0/1 ndarray.

    modality_type "location": (latitude, longitude) rows, haversine kNN, k = k_basis + 1;
    "time": (datetaken, dateupload) rows, |d taken| + |d upload|, k = 3 k_basis + 1;
    "text"/"cosine": rows are (TF-IDF or embedding) vectors, cosine top-(k_basis+1);
    "username": rows of the same user name are neighbours; "tags": Jaccard top-k_basis of the tag sets;
    anything else: Euclidean kNN with k = k_basis (the reference's default branch).
    """
    if modality_type in STRING_MODALITIES:
        W = len(_first_column(data))
        out = torch.zeros((W, W), dtype=torch.uint8, device=device)
        return _string_adjacency(data, modality_type, k_basis, out, device).cpu().numpy().astype(np.float64)
    A = np.asarray(data, dtype=np.float64)
    W = A.shape[0]
    out = np.zeros((W, W))
    valid, k, metric = _modality_plan(A, modality_type, k_basis)
    if len(valid) == 0:
        return out
    adj = topk_adjacency(A[valid], min(k, len(valid)), metric=metric, device=device).cpu().numpy()
    out[np.ix_(valid, valid)] = adj
    return out


def _modality_plan(A, modality_type, k_basis):
    """(valid row indices, k, metric) of one modality, matrix_operations.py:22-35, :101, :112-118."""
    if modality_type == "location":
        return np.where(~np.isnan(A).any(axis=1))[0], k_basis + 1, "haversine"
    if modality_type == "time":
        return np.where(~((A[:, 0] == 0.0) | (A[:, 1] == 0.0)))[0], 3 * k_basis + 1, "l1sum"
    if modality_type in ("text", "cosine"):
        return np.where(np.all(np.isfinite(A), axis=1))[0], k_basis + 1, "cosine"
    return np.where(np.all(np.isfinite(A), axis=1))[0], max(1, k_basis), "euclidean"


def fused_adjacency(data_modalities, modality_types, k_basis=50, *, device="cuda:0"):
    """create_adjacency_matrix of every modality (matrix_operations.py:14-132) OR-fused (:134-141) in ONE uint8
    W x W matrix that never leaves the device: every modality's top-k kernel ORs its edges straight into the
    accumulator (`mused_topk_adjacency(..., adj)` accumulates).  Returns a CUDA uint8 tensor."""
    W = len(data_modalities[0])
    fused = torch.zeros((W, W), dtype=torch.uint8, device=device)
    for m, t in zip(data_modalities, modality_types):
        if t in STRING_MODALITIES:
            _string_adjacency(m, t, k_basis, fused, device)
            continue
        A = np.asarray(m, dtype=np.float64)
        valid, k, metric = _modality_plan(A, t, k_basis)
        if len(valid) == 0:
            continue
        if len(valid) == W:
            topk_adjacency(A, min(k, W), metric=metric, out=fused, device=device)
        else:                                                          # rows with missing values are left out (:24, :35)
            sub = topk_adjacency(A[valid], min(k, len(valid)), metric=metric, device=device)
            idx = torch.as_tensor(valid, device=fused.device)
            fused[idx[:, None], idx[None, :]] |= sub
    return fused


def fuse_matrices(matrices, *, device="cuda:0"):
    """Element-wise OR of adjacency matrices -> int64 (matrix_operations.py:134-141).  Drop-in form for callers
    that hold per-modality matrices; the window loop (pipeline.py) uses fused_adjacency, which never
    materialises them."""
    acc = None
    for m in matrices:
        t = m if isinstance(m, torch.Tensor) else torch.as_tensor(np.asarray(m))
        t = t.to(device) if not t.is_cuda else t
        b = t != 0
        acc = b if acc is None else (acc | b)
    out = acc.to(torch.int64)
    if isinstance(matrices[0], torch.Tensor):
        return out
    return out.cpu().numpy().astype(int)


def perform_svd_reduction(matrix, reduced_dim, seed, *, n_iter=5, n_oversamples=10, device="cuda:0",
                          return_model=False):
    """matrix_operations.py:143-147: ``TruncatedSVD(n_components=min(reduced_dim, n_features - 1),
    random_state=seed).fit_transform(matrix)`` as a randomized SVD on the device (C ABI ``mused_svd_reduce``).

    ``matrix``: the fused adjacency, a CUDA uint8 tensor (``fused_adjacency``) or anything array-like (taken as
    float64).  The Gaussian test matrix is drawn here with numpy's ``RandomState(seed).normal`` exactly as sklearn's
    ``randomized_range_finder`` draws it, so the same seed starts from the same subspace; sklearn re-conditions the
    basis by LU, the device by Cholesky-QR -- the same range, results equal up to rounding
    (tests/test_svd_reduce_gpu.py).  Returns the W x k reduced matrix (numpy float64) [and singular values, V^T]."""
    lib = _lib.load()
    if isinstance(matrix, torch.Tensor) and matrix.dtype == torch.uint8:
        M, dt = matrix.to(device).contiguous(), _lib.U8
    else:
        src = matrix.detach().cpu().numpy() if isinstance(matrix, torch.Tensor) else np.asarray(matrix)
        M, dt = torch.as_tensor(np.ascontiguousarray(src, dtype=np.float64)).to(device), _lib.F64
    m, n = M.shape
    if m < n:
        raise ValueError("perform_svd_reduction on the device takes matrices with rows >= columns (got %d x %d)" % (m, n))
    k = min(int(reduced_dim), n - 1)
    q = min(k + int(n_oversamples), n)
    if q > 160:
        raise ValueError("n_components + n_oversamples = %d exceeds the device limit of 160" % q)
    omega = np.random.RandomState(seed).normal(size=(n, q)) if not isinstance(seed, np.random.RandomState) \
        else seed.normal(size=(n, q))
    Om = torch.as_tensor(omega).to(M.device)
    Xt = torch.empty((m, k), dtype=torch.float64, device=M.device)
    sig = torch.empty(k, dtype=torch.float64, device=M.device)
    vt = torch.empty((k, n), dtype=torch.float64, device=M.device)
    with torch.cuda.device(M.device):
        _lib.check(lib.mused_svd_reduce(M.data_ptr(), dt, m, n, M.stride(0), Om.data_ptr(), q, k, int(n_iter),
                                        Xt.data_ptr(), k, sig.data_ptr(), vt.data_ptr(), n,
                                        torch.cuda.current_stream().cuda_stream))
    out = Xt.cpu().numpy()
    return (out, sig.cpu().numpy(), vt.cpu().numpy()) if return_model else out


def contingency(prev_clusters, new_clusters, P, Q, *, device="cuda:0"):
    """counts[p, q] = #{i: prev[i] == p and new[i] == q} (matrix_operations.py:169-172)."""
    lib = _lib.load()
    p = torch.as_tensor(np.asarray(prev_clusters), dtype=torch.int32).to(device)
    q = torch.as_tensor(np.asarray(new_clusters), dtype=torch.int32).to(device)
    counts = torch.empty((P, Q), dtype=torch.int32, device=p.device)
    with torch.cuda.device(p.device):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.mused_contingency(p.data_ptr(), q.data_ptr(), p.numel(), P, Q, counts.data_ptr(), st))
    return counts


def match_clusters(prev_clusters, new_clusters, method="hungarian", min_overlap=5, *, device="cuda:0"):
    """matrix_operations.py:155-185 + :212-233: overlap counts on the GPU (one
    histogram pass instead of the O(P*Q*W) loop), Hungarian on the host."""
    from scipy.optimize import linear_sum_assignment
    if prev_clusters is None or len(prev_clusters) == 0:
        return new_clusters
    prev_clusters = np.asarray(prev_clusters)
    new_clusters = np.asarray(new_clusters)
    unique_prev, ip = np.unique(prev_clusters, return_inverse=True)
    unique_new, iq = np.unique(new_clusters, return_inverse=True)
    counts = contingency(ip, iq, len(unique_prev), len(unique_new), device=device).cpu().numpy()
    cost = np.where(counts >= min_overlap, -counts.astype(float), np.inf)
    if (np.all(np.isinf(cost)) or np.any(np.all(np.isinf(cost), axis=1))
            or np.any(np.all(np.isinf(cost), axis=0))):           # is_feasible :226-233
        return new_clusters
    if method != "hungarian":
        raise ValueError("Invalid method. Choose 'hungarian' or 'pot'.")  # 'pot' needs POT (absent)
    row_ind, col_ind = linear_sum_assignment(cost)
    mapping = {unique_new[c]: unique_prev[r] for r, c in zip(row_ind, col_ind)}
    return np.array([mapping.get(c, c) for c in new_clusters])


def nearest_centroid(X, centroids, *, device="cuda:0"):
    """argmin_j ||x_i - c_j|| on the GPU (KMeans E-step / predict; matrix_operations.py:286-289)."""
    return cosine_assign(X, centroids, -1.0, metric="l2", device=device)


def device_kmeans(matrix, n_clusters, seed=0, *, max_iter=300, tol=1e-4, device="cuda:0"):
    """KMeans entirely on the device (C ABI mused_kmeans): k-means++ seeding, Lloyd iterations on the fused
    assignment kernel, deterministic centre update.  Returns (labels int32 tensor, centroids float32 tensor,
    inertia, n_iter).  tol is sklearn's: relative to the mean feature variance (sklearn/cluster/_kmeans.py)."""
    lib = _lib.load()
    Xd = _dev_f32(matrix, device)
    n, d = Xd.shape
    mu = Xd.mean(dim=0, keepdim=True)
    Xc = (Xd - mu).contiguous()                       # distances are translation invariant; fp32 keeps more digits
    var = float(Xc.var(dim=0, unbiased=False).mean().item())
    cent = torch.empty((int(n_clusters), d), dtype=torch.float32, device=Xd.device)
    labels = torch.empty(n, dtype=torch.int32, device=Xd.device)
    inertia, n_iter = C.c_double(0.0), C.c_int32(0)
    with torch.cuda.device(Xd.device):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.mused_kmeans(Xc.data_ptr(), n, d, Xc.stride(0) if n > 1 else d, int(n_clusters), int(seed) & (2 ** 64 - 1),
                                    int(max_iter), float(tol) * var, cent.data_ptr(), labels.data_ptr(),
                                    C.byref(inertia), C.byref(n_iter), st))
    return labels, cent + mu, float(inertia.value), int(n_iter.value)


def perform_clustering(matrix, n_clusters, seed, *, device="cuda:0", return_model=False, engine="sklearn"):
    """KMeans labels of the reduced window (matrix_operations.py:149-153).

    ``KMeans(n_clusters, random_state=seed).fit`` stays sklearn's (k-means++ seeding and Lloyd
    updates are not reproducible bit for bit on another implementation, SURVEY.md 7.3-6, and label
    identity with the reference is the parity gate); the assignment of every item to its nearest
    centre -- what ``fit_predict`` returns -- is the device operator ``nearest_centroid``, run on rows and
    centres shifted by the centroid mean (distances are translation invariant; the fp32 kernel's
    ||x||^2 - 2 x.c + ||c||^2 loses digits to a large common offset).  Rows on which the fp32 device label
    differs from sklearn's float64 label (near-ties) keep sklearn's: the operator never fails where the
    reference would not; ``perform_clustering.last_disagreements`` counts them for the tests.
    """
    if engine == "device":
        # the throughput option: no sklearn fit, no host round trip of the W x l matrix (labels are KMeans labels of
        # another seeding: equal to sklearn's only up to the usual ambiguity; judged by inertia / NMI)
        lab, cent, inertia, n_iter = device_kmeans(matrix, n_clusters, seed, device=device)
        lab = lab.cpu().numpy().astype(np.int32)
        return (lab, cent) if return_model else lab
    from sklearn.cluster import KMeans
    M = np.ascontiguousarray(np.asarray(matrix, dtype=np.float64))
    km = KMeans(n_clusters=n_clusters, random_state=seed).fit(M)
    C = km.cluster_centers_
    mu = C.mean(axis=0, keepdims=True)
    lab = cosine_assign(M - mu, C - mu, -1.0, metric="l2", device=device)
    lab = lab.cpu().numpy().astype(km.labels_.dtype)
    diff = lab != km.labels_
    perform_clustering.last_disagreements = int(diff.sum())
    if diff.any():
        lab[diff] = km.labels_[diff]
    return (lab, km) if return_model else lab


perform_clustering.last_disagreements = 0

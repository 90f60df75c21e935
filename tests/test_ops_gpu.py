"""GPU parity of the similarity / assignment operators against the float64
NumPy oracle (oracle/ops_oracle.py).  Labels must be identical except rows
the oracle itself flags as near-ties (top-2 gap below the fp32 resolution)."""
import numpy as np
import pytest
import torch

from oracle import ops_oracle

pytestmark = pytest.mark.gpu
TIE = 2e-5     # documented near-tie margin for fp32 similarity keys


@pytest.mark.parametrize("n,d,k", [(1, 8, 1), (257, 96, 9), (2000, 512, 150), (1000, 1536, 150), (333, 100, 65)])
def test_cosine_assign(n, d, k):
    from mused_b200 import matrix_operations as mo
    from mused_b200 import synth
    rng = np.random.default_rng(n + d + k)
    C = rng.standard_normal((k, d)).astype(np.float32)
    X = (C[rng.integers(0, k, n)] + 0.8 * rng.standard_normal((n, d))).astype(np.float32)
    X[::17] = rng.standard_normal((len(X[::17]), d))
    tau = 0.5
    ref, best, gap = ops_oracle.cosine_assign(X, C, tau)
    lab, score = mo.cosine_assign(X, C, tau, return_score=True)
    lab, score = lab.cpu().numpy(), score.cpu().numpy()
    np.testing.assert_allclose(score, best, atol=2e-5)
    sure = (gap > TIE) & (np.abs(best - tau) > TIE)
    assert np.array_equal(lab[sure], ref[sure])
    assert sure.mean() > 0.99


def test_cosine_assign_tensor_core_tiles_and_simt_agree(monkeypatch):
    """n >= 148 * 256 takes the two-accumulator tile of assign_tc.cuh, a ragged tail included; the
    fp32 SIMT kernel (shapes the tensor-core path does not take) must give the same labels."""
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(11)
    n, d, k = 148 * 256 + 77, 256, 150
    C = rng.standard_normal((k, d)).astype(np.float32)
    X = (C[rng.integers(0, k, n)] + 0.8 * rng.standard_normal((n, d))).astype(np.float32)
    X[5] = 0.0
    ref, best, gap = ops_oracle.cosine_assign(X, C, 0.3)
    sure = (gap > TIE) & (np.abs(best - 0.3) > TIE)
    lab, score = mo.cosine_assign(X, C, 0.3, return_score=True)
    assert np.array_equal(lab.cpu().numpy()[sure], ref[sure]) and sure.mean() > 0.99
    np.testing.assert_allclose(score.cpu().numpy(), best, rtol=1e-4, atol=1e-4)
    l2ref, l2best, l2gap = ops_oracle.l2_assign(X, C, -1.0)
    l2 = mo.cosine_assign(X, C, -1.0, metric="l2").cpu().numpy()
    s2 = l2gap > 1e-3
    assert np.array_equal(l2[s2], l2ref[s2])
    monkeypatch.setenv("MUSED_ASSIGN_SIMT", "1")
    lab2 = mo.cosine_assign(X[:5000], C, 0.3).cpu().numpy()
    assert np.array_equal(lab2[sure[:5000]], ref[:5000][sure[:5000]])


def test_cosine_assign_edge_cases():
    from mused_b200 import matrix_operations as mo
    d = 40
    C = np.eye(3, d, dtype=np.float32)
    X = np.zeros((4, d), dtype=np.float32)
    X[1, 0] = 5
    X[2, 1] = -3           # anti-parallel to centroid 1: best cosine is 0 (centroids 0 and 2)
    X[3, :3] = 1           # exact three-way tie -> lowest index
    lab, sc = mo.cosine_assign(X, C, 0.0, return_score=True)
    assert lab.cpu().tolist() == [0, 0, 0, 0]
    lab = mo.cosine_assign(X, C, 0.5)
    assert lab.cpu().tolist() == [-1, 0, -1, 0]


@pytest.mark.parametrize("kind,n,d,k,tau", [("clustered", 40000, 512, 150, 0.5), ("noise", 30000, 1024, 150, 0.08),
                                            ("noise", 5000, 256, 3, -1.0), ("dups", 3000, 128, 40, 0.2)])
def test_cosine_assign_screening_path_gives_the_exact_labels(kind, n, d, k, tau):
    """Labels-only cosine calls take one TF32 screening pass + float32 re-scoring of the undecided rows
    (assign_tc.cuh).  Its labels must be the float64 oracle's outside the near-tie margin -- on clustered data
    (almost nothing re-scored), on pure noise with the threshold in the bulk of the best cosines (a large share
    re-scored, every pending type) and with duplicated centroids (exact ties -> lowest index)."""
    import ctypes as C_
    from mused_b200 import matrix_operations as mo, _lib
    rng = np.random.default_rng(k + d)
    C = rng.standard_normal((k, d)).astype(np.float32)
    if kind == "clustered":
        X = (C[rng.integers(0, k, n)] + 0.8 * rng.standard_normal((n, d))).astype(np.float32)
        X[::3] = rng.standard_normal((len(X[::3]), d))
    elif kind == "noise":
        X = rng.standard_normal((n, d)).astype(np.float32)
        X[7] = 0.0
    else:
        C[k // 2:] = C[:k - k // 2]                      # every centroid twice
        X = (C[rng.integers(0, k, n)] + 0.5 * rng.standard_normal((n, d))).astype(np.float32)
    ref, best, gap = ops_oracle.cosine_assign(X, C, tau)
    lib = _lib.load()
    p0, p1 = C_.c_uint64(0), C_.c_uint64(0)
    _lib.check(lib.mused_assign_pending(C_.byref(p0)))
    lab = mo.cosine_assign(X, C, tau).cpu().numpy()                       # screening + re-score
    _lib.check(lib.mused_assign_pending(C_.byref(p1)))
    lab3, _ = mo.cosine_assign(X, C, tau, return_score=True)              # 3xTF32
    lab3 = lab3.cpu().numpy()
    assert lab.min() >= -1 and lab.max() < k                              # no pending code survives
    sure = (gap > TIE) & (np.abs(best - tau) > TIE)
    if kind == "dups":
        assert np.array_equal(lab, lab3)                                  # exact ties resolved to the lowest index by both
        assert np.all(lab[lab >= 0] < k - k // 2)
    else:
        assert np.array_equal(lab[sure], ref[sure])
        assert np.array_equal(lab[sure], lab3[sure])
    frac = (p1.value - p0.value) / n
    print("%s: %.2f %% of the rows re-scored" % (kind, 100 * frac))
    if kind == "clustered":
        assert frac < 0.02


@pytest.mark.parametrize("n,d,k", [(1, 32, 1), (129, 32, 5), (777, 96, 17), (5000, 64, 256), (148 * 256 + 5, 128, 33),
                                   (2 * 148 * 256 + 131, 32, 150), (4097, 2048, 96)])
def test_cosine_assign_screening_shapes(n, d, k):
    """Shapes around the screening kernel's tiling: one K block, fewer K blocks than pipeline stages, one centroid,
    the widest accumulator (k = 256), both tile heights with ragged tails, several tiles per persistent CTA."""
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(n + d + k)
    C = rng.standard_normal((k, d)).astype(np.float32)
    X = (C[rng.integers(0, k, n)] + 0.7 * rng.standard_normal((n, d))).astype(np.float32)
    X[::5] = rng.standard_normal((len(X[::5]), d))
    tau = 0.35
    ref, best, gap = ops_oracle.cosine_assign(X, C, tau)
    lab = mo.cosine_assign(X, C, tau).cpu().numpy()
    assert lab.min() >= -1 and lab.max() < k
    sure = (gap > TIE) & (np.abs(best - tau) > TIE) if k > 1 else (np.abs(best - tau) > TIE)
    assert np.array_equal(lab[sure], ref[sure])


def test_screening_operand_semantics_of_the_tensor_core():
    """assign_tc.cuh's error bound assumes that kind::tf32 UMMAs use the upper 19 bits of the fp32 words TMA
    delivered (truncation, relative error < 2^-10) -- measured here.  Rows x = (1 + 2^-11 + 2^-13) * c with c = ones
    (TF32-exact): truncation screens cos = 1 / (1 + 2^-11 + 2^-13) = 0.99939, round-to-nearest would give 1.00037.
    With tau = 0.9983 and the bound 1.6e-3 the row is inside the threshold band (re-scored) only in the first case.
    Either way the label is exact (cos = 1 >= tau)."""
    import ctypes as C_
    from mused_b200 import matrix_operations as mo, _lib
    n, d = 512, 128
    Cn = np.ones((1, d), dtype=np.float32)
    X = np.full((n, d), 1.0 + 2.0 ** -11 + 2.0 ** -13, dtype=np.float32)
    lib = _lib.load()
    p0, p1 = C_.c_uint64(0), C_.c_uint64(0)
    _lib.check(lib.mused_assign_pending(C_.byref(p0)))
    lab = mo.cosine_assign(X, Cn, 0.9983).cpu().numpy()
    _lib.check(lib.mused_assign_pending(C_.byref(p1)))
    assert np.all(lab == 0)
    assert p1.value - p0.value == n, "the tensor core did not truncate the fp32 operand: revisit SCREEN_EPS"


@pytest.mark.parametrize("n,d,k", [(500, 50, 12), (64, 7, 3)])
def test_l2_assign(n, d, k):
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(5)
    C = rng.standard_normal((k, d)).astype(np.float32)
    X = rng.standard_normal((n, d)).astype(np.float32)
    ref, best, gap = ops_oracle.l2_assign(X, C, -1.0)
    lab, score = mo.cosine_assign(X, C, -1.0, metric="l2", return_score=True)
    sure = gap > 1e-4
    assert np.array_equal(lab.cpu().numpy()[sure], ref[sure])
    np.testing.assert_allclose(score.cpu().numpy(), best, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("metric", ["cosine", "euclidean"])
@pytest.mark.parametrize("W,f,k", [(300, 24, 11), (2000, 64, 51), (5, 3, 9)])
def test_topk_adjacency(W, f, k, metric):
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(W + f)
    F = rng.standard_normal((W, f)).astype(np.float32)
    ref, margin = ops_oracle.knn_adjacency(F, k, metric)
    adj = mo.topk_adjacency(F, k, metric=metric).cpu().numpy()
    sure = margin > (1e-4 if metric == "euclidean" else TIE)
    assert np.array_equal(adj[sure], ref[sure])
    assert sure.mean() > 0.97 or W < 10
    # rows the oracle flags as near-ties still have exactly k (or k-1 without self) edges
    assert np.all(adj.sum(1) <= min(k, W)) and np.all(adj.diagonal() == 0)
    # OR-accumulation (fuse_matrices semantics): accumulating twice changes nothing,
    # accumulating the transpose-like second modality only adds edges
    out = torch.as_tensor(adj.copy()).cuda()
    mo.topk_adjacency(F, k, metric=metric, out=out)
    assert np.array_equal(out.cpu().numpy(), adj)


@pytest.mark.parametrize("W,k", [(400, 51), (2000, 151), (3, 7)])
def test_topk_adjacency_location_and_time(W, k):
    """float64 modalities of create_adjacency_matrix (matrix_operations.py:22-53)."""
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(W)
    loc = np.column_stack([rng.uniform(-70, 70, W), rng.uniform(-180, 180, W)])
    ref, margin = ops_oracle.knn_adjacency(loc, k, "haversine")
    adj = mo.topk_adjacency(loc, k, metric="haversine").cpu().numpy()
    sure = margin > 1e-9
    assert np.array_equal(adj[sure], ref[sure]) and (sure.mean() > 0.99 or W < 10)
    tm = np.column_stack([1.3e9 + rng.uniform(0, 3e6, W), 1.3e9 + rng.uniform(0, 3e6, W)])
    ref, margin = ops_oracle.knn_adjacency(tm, k, "l1sum")
    adj = mo.topk_adjacency(tm, k, metric="l1sum").cpu().numpy()
    sure = margin > 0                                    # same float64 operations: exact unless tied
    assert np.array_equal(adj[sure], ref[sure]) and (sure.mean() > 0.99 or W < 10)
    assert np.all(adj.diagonal() == 0)


def test_create_adjacency_location_time_validity_rules():
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(2)
    loc = np.column_stack([rng.uniform(-60, 60, 90), rng.uniform(-180, 180, 90)])
    loc[11, 1] = np.nan
    A = mo.create_adjacency_matrix(loc, "location", k_basis=6)
    valid = np.where(~np.isnan(loc).any(axis=1))[0]
    ref, margin = ops_oracle.knn_adjacency(loc[valid], 7, "haversine")
    assert np.array_equal(A[np.ix_(valid, valid)], ref.astype(float))
    assert A[11].sum() == 0 and A[:, 11].sum() == 0
    tm = np.column_stack([1.3e9 + rng.uniform(0, 1e6, 90), 1.3e9 + rng.uniform(0, 1e6, 90)])
    tm[3, 1] = 0.0
    A = mo.create_adjacency_matrix(tm, "time", k_basis=4)
    valid = np.where(~((tm[:, 0] == 0.0) | (tm[:, 1] == 0.0)))[0]
    ref, margin = ops_oracle.knn_adjacency(tm[valid], 13, "l1sum")
    assert np.array_equal(A[np.ix_(valid, valid)], ref.astype(float))
    assert A[3].sum() == 0 and A[:, 3].sum() == 0


def test_create_adjacency_and_fuse_match_reference_semantics():
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(9)
    F = rng.standard_normal((120, 6))
    F[7] = np.nan                                   # invalid rows are skipped (matrix_operations.py:114)
    A = mo.create_adjacency_matrix(F, "dense", k_basis=5)
    valid = np.where(np.all(np.isfinite(F), axis=1))[0]
    ref, margin = ops_oracle.knn_adjacency(F[valid], 5, "euclidean")
    sure = margin > 1e-4
    assert np.array_equal(A[np.ix_(valid, valid)][sure], ref[sure].astype(float))
    assert A[7].sum() == 0 and A[:, 7].sum() == 0 and A.dtype == np.float64
    fused = mo.fuse_matrices([A, A.T])
    assert fused.dtype.kind == "i" and np.array_equal(fused, ((A + A.T) > 0).astype(int))


def test_contingency_and_match_clusters():
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(4)
    prev = rng.integers(0, 7, 5000)
    cur = (prev + 3) % 7
    cur[rng.random(5000) < 0.05] = 6
    cnt = mo.contingency(prev, cur, 7, 7).cpu().numpy()
    assert np.array_equal(cnt, ops_oracle.contingency(prev, cur, 7, 7))
    assert cnt.sum() == 5000
    out = mo.match_clusters(prev, cur, method="hungarian", min_overlap=3)
    assert (out == prev).mean() > 0.9
    assert mo.match_clusters(None, cur) is cur


def test_device_kmeans_blobs_inertia_and_determinism():
    """mused_kmeans (perform_clustering's device engine, matrix_operations.py:149-153): exact recovery of
    well-separated blobs, inertia within 5 % of sklearn's KMeans on overlapping data, identical results for
    identical (X, k, seed)."""
    from sklearn.cluster import KMeans
    from sklearn.metrics import adjusted_rand_score
    from mused_b200 import matrix_operations as mo
    rng = np.random.default_rng(4)
    k, d, per = 12, 50, 160
    centres = rng.standard_normal((k, d)) * 6.0
    X = np.concatenate([centres[j] + 0.3 * rng.standard_normal((per, d)) for j in range(k)]) + 40.0   # far from the origin
    truth = np.repeat(np.arange(k), per)
    lab, cent, inertia, n_iter = mo.device_kmeans(X, k, seed=3)
    lab = lab.cpu().numpy()
    km = KMeans(n_clusters=k, random_state=0, n_init=10).fit(X)    # sklearn's best of 10 seedings
    assert inertia <= 1.05 * km.inertia_ or adjusted_rand_score(truth, lab) == 1.0
    lab2, cent2, inertia2, _ = mo.device_kmeans(X, k, seed=3)
    assert np.array_equal(lab, lab2.cpu().numpy()) and inertia == inertia2
    # labels are the nearest final centre (what fit_predict returns)
    D = ((X[:, None, :] - cent.cpu().numpy().astype(np.float64)[None, :, :]) ** 2).sum(-1)
    srt = np.sort(D, axis=1)
    sure = srt[:, 1] - srt[:, 0] > 1e-3 * srt[:, 0]
    assert np.array_equal(lab[sure], np.argmin(D, axis=1)[sure])
    assert abs(inertia - D.min(axis=1).sum()) <= 1e-3 * inertia
    # overlapping clusters: local optima differ between seedings; compare quality, not labels
    Y = rng.standard_normal((3000, 40)) + rng.standard_normal((15, 40))[rng.integers(0, 15, 3000)] * 1.5
    best = min(mo.device_kmeans(Y, 15, seed=s)[2] for s in range(3))
    ref = KMeans(n_clusters=15, random_state=0).fit(Y).inertia_
    assert best <= 1.05 * ref, (best, ref)
    out = mo.perform_clustering(Y, 15, 0, engine="device")
    assert out.shape == (3000,) and out.min() >= 0 and out.max() < 15

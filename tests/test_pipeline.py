"""SWFDMC window loop (main.py:25-119): oracle on CPU and device operators on the GPU against the
reference's own outputs (tests/golden/pipeline_swfdmc.npz, made by oracle/gen_pipeline_golden.py)."""
import os

import numpy as np
import pytest

from oracle import ops_oracle, swfd_oracle as so

GOLD = os.path.join(os.path.dirname(__file__), "golden", "pipeline_swfdmc.npz")
TYPES = ["location", "time", "visual"]


def _valid(data, t):
    if t == "location":
        return np.where(~np.isnan(data).any(axis=1))[0], "haversine", lambda kb: kb + 1
    if t == "time":
        return np.where(~((data[:, 0] == 0.0) | (data[:, 1] == 0.0)))[0], "l1sum", lambda kb: 3 * kb + 1
    return np.where(np.all(np.isfinite(data), axis=1))[0], "euclidean", lambda kb: max(1, kb)


def _oracle_adjacency(data, t, kb):
    W = len(data)
    out = np.zeros((W, W), dtype=np.uint8)
    valid, metric, kf = _valid(data, t)
    adj, margin = ops_oracle.knn_adjacency(data[valid], min(kf(kb), len(valid)), metric)
    out[np.ix_(valid, valid)] = adj
    return out, margin


def test_oracle_window_loop_matches_reference_golden():
    g = np.load(GOLD, allow_pickle=False)
    W, NWIN, REDUCED, KB, SEED = (int(v) for v in g["params"])
    mods = [g["loc"], g["tim"], g["vis"]]
    sk = None
    for w in range(NWIN):
        lo, hi = w * W, (w + 1) * W
        fused = np.zeros((W, W), dtype=np.uint8)
        for m, t in zip(mods, TYPES):
            adj, margin = _oracle_adjacency(m[lo:hi], t, KB)
            assert margin.min() > 0                         # no ties in the fixture: the comparison is exact
            fused |= adj
        assert np.array_equal(fused, g["fused"][w])
        F = fused.astype(np.int64)
        if sk is None:
            sk = so.SeqSWFD(W, float(np.max(np.linalg.norm(F, axis=1) ** 2)), W, REDUCED)
        for i in range(W):
            sk.fit(F[i:i + 1])
        red = sk.get()[0]
        red = red if red.shape[0] == W else red.T
        assert np.array_equal(red, g["reduced"][w])        # the restatement is bit-identical to the reference


@pytest.mark.gpu
def test_device_window_loop_labels_and_scores_match_reference_golden():
    from sklearn.metrics import f1_score, normalized_mutual_info_score
    from mused_b200 import pipeline
    g = np.load(GOLD, allow_pickle=False)
    W, NWIN, REDUCED, KB, SEED = (int(v) for v in g["params"])
    mods = [g["loc"], g["tim"], g["vis"]]
    clusters, true, reduced = pipeline.process_streaming_data(
        mods, TYPES, g["labels"], window_size=W, reduced_dim=REDUCED, k_basis=KB, seed=SEED, schedule="reference")
    assert np.array_equal(true, g["labels"])
    for w in range(NWIN):
        a, b = reduced[w], g["reduced"][w]
        sign = np.sign(np.einsum("ij,ij->j", a, b))
        sign[sign == 0] = 1.0
        np.testing.assert_allclose(a * sign[None, :], b, rtol=0, atol=1e-7 * np.abs(b).max())
    assert np.array_equal(clusters, g["clusters"])          # identical event assignments
    assert normalized_mutual_info_score(true, clusters) == pytest.approx(float(g["nmi"]), abs=1e-12)
    assert f1_score(true, clusters, average="weighted", zero_division=0) == pytest.approx(float(g["f1"]), abs=1e-12)
    ev = true > 0
    assert normalized_mutual_info_score(true[ev], clusters[ev]) == pytest.approx(float(g["nmi_e"]), abs=1e-12)


@pytest.mark.gpu
def test_configs2_shape_window_loop_matches_reference_loop():
    """BASELINE configs[2] shape (W = 2000, reduced_dim = 50, k_basis = 50, location + time + two dense
    embeddings of 512 + 768 columns): the device loop with schedule="reference" returns the labels of the
    UNMODIFIED reference loop (oracle/time_reference_c3.py -> tests/golden/c3_small.npz) on two windows, hence the
    same NMI / F1; schedule="batched" with the device KMeans is scored, not label-compared."""
    from sklearn.metrics import f1_score, normalized_mutual_info_score
    from mused_b200 import pipeline, synth
    p = os.path.join(os.path.dirname(__file__), "golden", "c3_small.npz")
    if not os.path.exists(p):
        pytest.skip("c3_small.npz missing")
    g = np.load(p)
    n, W = int(g["n_items"]), int(g["W"])
    mods, _, labels = synth.multimodal_event_stream(n, n_events=150, noise_rate=0.95)
    types = ["location", "time", "visual", "embedding"]
    m = int(g["n_windows"]) * W
    clusters, true, _ = pipeline.process_streaming_data([x[:m] for x in mods], types, labels[:m], window_size=W,
                                                        reduced_dim=int(g["reduced_dim"]), k_basis=int(g["k_basis"]),
                                                        schedule="reference")
    same = int(np.sum(clusters == g["clusters"]))
    assert same >= m - 2, "%d of %d labels differ from the reference loop" % (m - same, m)   # kNN / KMeans near-ties
    if same == m:
        assert normalized_mutual_info_score(true, clusters) == pytest.approx(float(g["nmi"]), abs=1e-12)
        assert f1_score(true, clusters, average="weighted", zero_division=0) == pytest.approx(float(g["f1"]), abs=1e-12)
    c2, t2, _ = pipeline.process_streaming_data([x[:m] for x in mods], types, labels[:m], window_size=W, reduced_dim=50,
                                                k_basis=50, schedule="batched", kmeans="device")
    assert len(c2) == m and normalized_mutual_info_score(t2, c2) > 0.5 * float(g["nmi"])

"""perform_svd_reduction on the device (C ABI mused_svd_reduce: randomized TruncatedSVD of the fused adjacency,
matrix_operations.py:143-147) against

  * tests/golden/svd_reduce.npz -- outputs of the UNMODIFIED reference function on seeded community-structured 0/1
    matrices (oracle/gen_golden_svd.py), and
  * sklearn's TruncatedSVD run live on a window-sized kNN adjacency built by the device top-k kernels.

The reduced matrix X V is compared through what the clustering step sees: singular values (column norms), the
Gram X_t X_t^T (free of sign and of rotations inside near-equal singular values) and, where the spectrum has gaps,
the columns themselves (svd_flip's sign rule is reproduced).
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _check(Xd, Xr, tag):
    assert Xd.shape == Xr.shape, tag
    sd, sr = np.linalg.norm(Xd, axis=0), np.linalg.norm(Xr, axis=0)
    np.testing.assert_allclose(sd, sr, rtol=1e-7, atol=1e-9 * sr.max(), err_msg=tag)
    Gd, Gr = Xd @ Xd.T, Xr @ Xr.T
    assert np.abs(Gd - Gr).max() <= 1e-7 * np.abs(Gr).max(), tag
    # columns whose singular value is separated from its neighbours by > 1 %: equal including the sign
    gap = np.minimum(np.abs(np.diff(sr, prepend=np.inf)), np.abs(np.diff(sr, append=0.0))) / sr
    lead = gap > 1e-2
    assert lead[:2].all(), tag
    assert np.abs(Xd[:, lead] - Xr[:, lead]).max() <= 1e-6 * sr.max(), tag


def test_svd_reduce_matches_reference_fixture(golden_dir):
    from oracle.gen_golden_svd import block_adjacency
    from mused_b200 import matrix_operations as mo
    p = os.path.join(golden_dir, "svd_reduce.npz")
    if not os.path.exists(p):
        pytest.skip("fixture missing")
    g = np.load(p)
    for i in range(int(g["n_cases"])):
        W, blocks, seed, reduced_dim, rs = (int(v) for v in g["case%d_args" % i])
        p_in, p_out = (float(v) for v in g["case%d_p" % i])
        M = block_adjacency(W, blocks, p_in, p_out, seed)
        assert int(M.sum()) == int(g["case%d_sum" % i]), "generator drifted from the fixture"
        Xr = g["case%d_X" % i]
        Xd = mo.perform_svd_reduction(torch.as_tensor(M).cuda(), reduced_dim, rs)          # uint8 device matrix
        _check(Xd, Xr, "case %d (uint8)" % i)
        Xf = mo.perform_svd_reduction(M.astype(np.int64), reduced_dim, rs)                  # host int matrix -> float64 path
        _check(Xf, Xr, "case %d (float64)" % i)


def test_svd_reduce_matches_sklearn_on_a_window_adjacency():
    from sklearn.decomposition import TruncatedSVD
    from mused_b200 import matrix_operations as mo, synth
    A, labels, centres = synth.event_stream(2000, blocks=(256, 192, 32, 32), n_events=150, seed=5)
    fused = mo.fused_adjacency([A[:, :256], A[:, 256:448]], ["dense", "dense"], k_basis=50)
    M = fused.cpu().numpy().astype(int)
    Xr = TruncatedSVD(n_components=50, random_state=0).fit_transform(M)                     # = matrix_operations.py:143-147
    Xd, sig, vt = mo.perform_svd_reduction(fused, 50, 0, return_model=True)
    _check(Xd, Xr, "window adjacency")
    assert np.abs(vt @ vt.T - np.eye(50)).max() < 1e-10                                      # orthonormal right vectors
    assert np.all(np.diff(sig) <= 1e-12 * sig[0])
    assert np.all(np.abs(vt[np.arange(50), np.abs(vt).argmax(axis=1)] - np.abs(vt).max(axis=1)) < 1e-15)   # svd_flip rule


def test_ssvdmc_pipeline_matches_host_sklearn_loop():
    """approach="sSVDMC" (main.py:77-79 + :97 + :110): device adjacency + device TruncatedSVD + KMeans + matching
    against the same loop with sklearn's TruncatedSVD on the host -- identical labels."""
    from sklearn.decomposition import TruncatedSVD
    from mused_b200 import matrix_operations as mo, pipeline, synth
    A, labels, _ = synth.event_stream(1200, blocks=(64, 48), n_events=12, seed=9)
    mods, types = [A[:, :64], A[:, 64:]], ["dense", "dense"]
    W = 400
    got, true, red = pipeline.process_streaming_data(mods, types, labels, window_size=W, reduced_dim=20, k_basis=15,
                                                     seed=0, approach="sSVDMC")
    prev, want = None, []
    for lo in range(0, 1200, W):
        fused = mo.fused_adjacency([m[lo:lo + W] for m in mods], types, k_basis=15).cpu().numpy().astype(int)
        r = TruncatedSVD(n_components=20, random_state=0).fit_transform(fused)
        c = mo.perform_clustering(r, len(np.unique(labels[lo:lo + W])), 0)
        c = mo.match_clusters(prev, c, method="hungarian", min_overlap=3)
        prev = c
        want.extend(c)
    assert np.array_equal(got, np.asarray(want))

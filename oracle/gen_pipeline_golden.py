"""Golden vectors for the SWFDMC window loop (main.py:25-119) from the UNMODIFIED reference.

Runs the reference's own functions -- create_adjacency_matrix, fuse_matrices,
swfd.SeqBasedSWFD, perform_clustering, match_clusters, metrics_evaluation.compute_all_metrics --
in the order of main.process_streaming_data(approach="SWFDMC") on a seeded synthetic multimodal
event stream (location + time + visual features) and stores inputs and outputs in
tests/golden/pipeline_swfdmc.npz.  Build container only:  python -m oracle.gen_pipeline_golden
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import contextlib
import io
import os

import numpy as np
import scipy
import sklearn

from . import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "pipeline_swfdmc.npz")
W, NWIN, REDUCED, KB, SEED = 240, 3, 12, 8, 0


def make_stream(n, n_events=6, noise_rate=0.4, seed=0):
    rng = np.random.default_rng([seed, 41])
    ev_loc = np.column_stack([rng.uniform(-50, 50, n_events), rng.uniform(-150, 150, n_events)])
    ev_time = 1.30e9 + rng.uniform(0, 5e6, n_events)
    ev_vis = rng.standard_normal((n_events, 16)) * 3
    labels = np.zeros(n, dtype=np.int64)
    loc = np.empty((n, 2)); tim = np.empty((n, 2)); vis = np.empty((n, 16))
    for i in range(n):
        active = (i // 80) % n_events                      # events arrive in bursts
        if rng.random() < noise_rate:
            loc[i] = [rng.uniform(-60, 60), rng.uniform(-180, 180)]
            t = 1.30e9 + rng.uniform(0, 5e6)
            tim[i] = [t, t + rng.uniform(0, 3e5)]
            vis[i] = rng.standard_normal(16) * 3
        else:
            e = active if rng.random() < 0.7 else int(rng.integers(0, n_events))
            labels[i] = e + 1
            loc[i] = ev_loc[e] + rng.normal(0, 0.05, 2)
            t = ev_time[e] + rng.normal(0, 2000)
            tim[i] = [t, t + abs(rng.normal(0, 5000))]
            vis[i] = ev_vis[e] + rng.standard_normal(16) * 0.3
    loc[7] = np.nan                                        # invalid rows (matrix_operations.py:24, :35)
    tim[11, 0] = 0.0
    return loc, tim, vis, labels


def main():
    swfd_pkg, mo, me = ref_shim.load()
    n = W * NWIN
    loc, tim, vis, labels = make_stream(n, seed=SEED)
    mods, types = [loc, tim, vis], ["location", "time", "visual"]
    sk, prev = None, None
    all_clusters, fused_all, red_all = [], [], []
    sink = io.StringIO()
    for w in range(NWIN):
        lo, hi = w * W, (w + 1) * W
        true = labels[lo:hi]
        n_clusters = len(np.unique(true))
        with contextlib.redirect_stdout(sink):
            mats = [mo.create_adjacency_matrix(data=m[lo:hi], modality_type=t, k_basis=KB) for m, t in zip(mods, types)]
        fused = mo.fuse_matrices(mats)
        if sk is None:
            max_norm = np.max(np.linalg.norm(fused, axis=1) ** 2)
            sk = swfd_pkg.SeqBasedSWFD(N=W, R=max_norm, d=fused.shape[1], sketch_dim=REDUCED)
        for i in range(fused.shape[0]):
            sk.fit(fused[i, :].reshape(1, -1))
        red, _, _, _ = sk.get()
        if red.shape[0] != W:
            red = red.T
        clusters = mo.perform_clustering(red, n_clusters, SEED)
        clusters = mo.match_clusters(prev, clusters, method="hungarian", min_overlap=3)
        prev = clusters
        all_clusters.extend(clusters)
        fused_all.append(fused.astype(np.uint8))
        red_all.append(np.array(red))
    results = {"subset_size": [], "noise_rate": [], "label_mode": [], "sorting": [], "reduced_dim": [], "k_basis": [],
               "window_size": [], "nmi_score": [], "nmi_e_score": [], "f1_score": []}
    with contextlib.redirect_stdout(sink):
        me.compute_all_metrics(results, n, 0.4, "x", "x", REDUCED, KB, W, list(all_clusters), list(labels), 1, 0)
    np.savez_compressed(
        OUT, loc=loc, tim=tim, vis=vis, labels=labels, clusters=np.asarray(all_clusters), fused=np.stack(fused_all),
        reduced=np.stack(red_all), nmi=results["nmi_score"][0], nmi_e=results["nmi_e_score"][0],
        f1=results["f1_score"][0], params=np.array([W, NWIN, REDUCED, KB, SEED]),
        versions="numpy %s scipy %s sklearn %s" % (np.__version__, scipy.__version__, sklearn.__version__))
    print("wrote", OUT, "nmi %.4f nmi_e %.4f f1 %.4f" % (results["nmi_score"][0], results["nmi_e_score"][0],
                                                         results["f1_score"][0]))


if __name__ == "__main__":
    main()

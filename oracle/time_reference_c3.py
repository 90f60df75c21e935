"""BASELINE configs[2] (MediaEval-SED-shaped stream, W = 2000, reduced_dim = 50, k_basis = 50; main.py:309-311):
the UNMODIFIED reference's own SWFDMC window loop (main.py:25-119: create_adjacency_matrix, fuse_matrices,
SeqBasedSWFD.fit row by row, get, perform_clustering, match_clusters) on the first windows of the synthetic
stream of mused_b200.synth.multimodal_event_stream, timed, with its labels and scores stored in
tests/golden/c3_prefix.npz for the GPU loop to be compared with (tools/run_c3.py).
Build container only:  OMP_NUM_THREADS=1 python -m oracle.time_reference_c3 [n_windows=2]
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import contextlib
import io
import os
import sys
import time

import numpy as np

from . import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth  # noqa: E402

W, REDUCED, KB, SEED, N_ITEMS = 2000, 50, 50, 0, 450000
TYPES = ["location", "time", "visual", "embedding"]        # the two dense modalities take the reference's default branch


def main():
    nwin = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    n_items = int(sys.argv[2]) if len(sys.argv) > 2 else N_ITEMS      # the generator's draws depend on the stream length
    swfd_pkg, mo, me = ref_shim.load()
    mods, _, labels = synth.multimodal_event_stream(n_items, n_events=150, noise_rate=0.95)
    sk, prev = None, None
    all_clusters = []
    sink = io.StringIO()
    t0 = time.perf_counter()
    parts = {"adjacency": 0.0, "sketch": 0.0, "cluster": 0.0}
    for w in range(nwin):
        lo, hi = w * W, (w + 1) * W
        true = labels[lo:hi]
        n_clusters = len(np.unique(true))
        ta = time.perf_counter()
        with contextlib.redirect_stdout(sink):
            mats = [mo.create_adjacency_matrix(data=m[lo:hi], modality_type=t, k_basis=KB) for m, t in zip(mods, TYPES)]
        fused = mo.fuse_matrices(mats)
        tb = time.perf_counter()
        if sk is None:
            max_norm = np.max(np.linalg.norm(fused, axis=1) ** 2)
            sk = swfd_pkg.SeqBasedSWFD(N=W, R=max_norm, d=fused.shape[1], sketch_dim=REDUCED)
        for i in range(fused.shape[0]):
            sk.fit(fused[i, :].reshape(1, -1))
        red, _, _, _ = sk.get()
        if red.shape[0] != W:
            red = red.T
        tc = time.perf_counter()
        clusters = mo.perform_clustering(red, n_clusters, SEED)
        clusters = mo.match_clusters(prev, clusters, method="hungarian", min_overlap=3)
        td = time.perf_counter()
        prev = clusters
        all_clusters.extend(clusters)
        parts["adjacency"] += tb - ta
        parts["sketch"] += tc - tb
        parts["cluster"] += td - tc
        print("window %d: adjacency %.1fs sketch %.1fs cluster %.1fs (levels %d)" % (w, tb - ta, tc - tb, td - tc, sk.logR), flush=True)
    dt = time.perf_counter() - t0
    n = nwin * W
    res, _ = me.get_initial_results()
    with contextlib.redirect_stdout(sink):
        me.compute_all_metrics(res, n, 0.95, "multi", "none", REDUCED, KB, W, list(all_clusters), list(labels[:n]), 1, 0)
    out = os.path.join(ROOT, "tests", "golden", "c3_prefix.npz" if n_items == N_ITEMS else "c3_small.npz")
    np.savez_compressed(out, n_items=n_items, clusters=np.asarray(all_clusters), n_windows=nwin, W=W, reduced_dim=REDUCED, k_basis=KB,
                        seconds=dt, items_per_s=n / dt, nmi=res["nmi_score"][0], nmi_e=res["nmi_e_score"][0],
                        f1=res["f1_score"][0], parts=np.array([parts["adjacency"], parts["sketch"], parts["cluster"]]))
    print("reference loop: %d items in %.1fs -> %.1f items/s (adjacency %.1fs, sketch %.1fs, cluster+match %.1fs); "
          "NMI %.4f NMI_e %.4f F1 %.4f; wrote %s" % (n, dt, n / dt, parts["adjacency"], parts["sketch"], parts["cluster"],
                                                   res["nmi_score"][0], res["nmi_e_score"][0], res["f1_score"][0], out))


if __name__ == "__main__":
    main()

"""BASELINE configs[2] at size, end to end on the device operators: ~450k MediaEval-SED-shaped items (location +
time + visual + text embeddings, 1536 feature columns), windows of W = 2000 (main.py:309), per window kNN adjacency
per modality -> one fused uint8 matrix on the device -> SWFD (l = 50, d = W) -> KMeans -> Hungarian matching
(main.py:25-119); NMI / F1 with the same sklearn calls as metrics_evaluation.py:48-72.  The first windows are
compared with the labels of the UNMODIFIED reference loop (tests/golden/c3_prefix.npz, oracle/time_reference_c3.py).
Usage: python tools/run_c3.py [n_items=450000] [schedules=reference,batched]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sklearn.metrics import f1_score, normalized_mutual_info_score   # noqa: E402
from mused_b200 import pipeline, synth                                # noqa: E402

n_items = int(sys.argv[1]) if len(sys.argv) > 1 else 450000
schedules = sys.argv[2].split(",") if len(sys.argv) > 2 else ["reference", "batched"]
W = 2000
TYPES = ["location", "time", "visual", "embedding"]       # dense modalities: the reference's default (Euclidean) branch
n_items -= n_items % W
mods, _, labels = synth.multimodal_event_stream(450000, n_events=150, noise_rate=0.95)
mods = [m[:n_items] for m in mods]
labels = labels[:n_items]
fx = os.path.join(ROOT, "tests", "golden", "c3_prefix.npz")
g = np.load(fx) if os.path.exists(fx) else None
for schedule in schedules:
    pipeline.process_streaming_data([m[:W] for m in mods], TYPES, labels[:W], window_size=W, reduced_dim=50, k_basis=50,
                                    schedule=schedule)  # warm-up
    t0 = time.perf_counter()
    clusters, true, _ = pipeline.process_streaming_data(mods, TYPES, labels, window_size=W, reduced_dim=50, k_basis=50,
                                                        schedule=schedule)
    dt = time.perf_counter() - t0
    ev = true > 0
    line = ("configs[2], schedule=%s: %d items in %d windows of %d: %.1f s -> %.0f items/s; NMI %.4f NMI_e %.4f F1(weighted) %.4f"
            % (schedule, len(true), len(true) // W, W, dt, len(true) / dt, normalized_mutual_info_score(true, clusters),
               normalized_mutual_info_score(true[ev], clusters[ev]) if ev.sum() > 1 else 0.0,
               f1_score(true, clusters, average="weighted", zero_division=0)))
    if g is not None:
        m = int(g["n_windows"]) * W
        same = int(np.sum(clusters[:m] == g["clusters"]))
        line += ("; first %d items vs the reference's own loop (%.1f items/s on the build container's CPU): %d / %d labels identical"
                 % (m, float(g["items_per_s"]), same, m))
    print(line, flush=True)

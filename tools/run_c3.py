"""BASELINE configs[2] shape end to end on the device operators: MediaEval-SED-like multimodal items
(location + time + visual + text embeddings, d = 1536 of features), windows of W = 2000 items
(main.py:309), kNN adjacency per modality -> fusion -> SWFD (l = 50, d = W) -> KMeans -> Hungarian
matching -> NMI / F1 (main.py:25-119).  Prints items/s and the scores."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from sklearn.metrics import f1_score, normalized_mutual_info_score
from mused_b200 import pipeline, synth

nwin = int(sys.argv[1]) if len(sys.argv) > 1 else 5
W = 2000
mods, types, labels = synth.multimodal_event_stream(nwin * W, n_events=150, noise_rate=float(sys.argv[2]) if len(sys.argv) > 2 else 0.95)
for schedule in ("reference", "batched"):
    pipeline.process_streaming_data([m[:W] for m in mods], types, labels[:W], window_size=W, reduced_dim=50, k_basis=50,
                                    schedule=schedule)  # warm-up
    t0 = time.perf_counter()
    clusters, true, _ = pipeline.process_streaming_data(mods, types, labels, window_size=W, reduced_dim=50, k_basis=50,
                                                        schedule=schedule)
    dt = time.perf_counter() - t0
    ev = true > 0
    print("configs[2] shape, schedule=%s: %d items in %d windows of %d: %.1f s -> %.0f items/s; NMI %.4f NMI_e %.4f F1(weighted) %.4f" %
          (schedule, len(true), nwin, W, dt, len(true) / dt, normalized_mutual_info_score(true, clusters),
           normalized_mutual_info_score(true[ev], clusters[ev]) if ev.sum() > 1 else 0.0,
           f1_score(true, clusters, average="weighted", zero_division=0)), flush=True)

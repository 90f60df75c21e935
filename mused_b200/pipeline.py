"""The SWFDMC window loop of /root/reference/main.py:25-119 on the device operators (and, with
``approach="sSVDMC"``, the reference's per-window TruncatedSVD baseline, main.py:77-79, on the same operators).

Same steps, same order, same arguments as ``process_streaming_data(approach="SWFDMC")``:
per full window -- one kNN adjacency per modality (create_adjacency_matrix, :44-49), OR-fusion
(:56), the rows of the fused W x W matrix through the sliding-window FD sketch (:60-67), ``get()``
and the transpose fix-up (:70-76), KMeans (:97), Hungarian matching with the previous window
(:106-110).  The fused adjacency stays on the device as one uint8 matrix from the top-k kernels to the sketch's
row staging kernel (no per-modality W x W float64 matrices, no host round trip).  The "username" and "tags"
modalities are taken as string columns (ids on the host, pairwise work on the device); "text" takes the TF-IDF (or
embedding) matrix the caller built.
"""
from __future__ import annotations

import numpy as np
import torch

from . import matrix_operations as mo
from .swfd import SeqBasedSWFD


def process_streaming_data(data_modalities, modality_types, complete_true_labels, *, window_size=2000,
                           reduced_dim=50, k_basis=50, seed=0, step_window_ratio=1, schedule="reference",
                           device="cuda:0", kmeans="sklearn", approach="SWFDMC"):
    """Returns (all_clusters, all_true_labels, per-window list of reduced matrices)."""
    if approach not in ("SWFDMC", "sSVDMC"):
        raise ValueError("approach must be SWFDMC or sSVDMC (the other baselines of main.py are host sklearn code)")
    subset_size = len(complete_true_labels)
    window, all_clusters, all_true, reduced = [], [], [], []
    swfd, prev_clusters = None, None
    for i in range(subset_size):
        window.append(i)
        if len(window) > window_size:
            window.pop(0)
        if len(window) == window_size and (i + 1) * step_window_ratio % window_size == 0:   # main.py:32
            lo = i - window_size + 1
            true_labels = np.asarray(complete_true_labels[lo:i + 1])
            all_true.extend(true_labels)
            n_clusters = len(np.unique(true_labels))                                          # main.py:41
            # main.py:44-56: one uint8 W x W matrix on the device, every modality's top-k kernel ORs into it
            fused = mo.fused_adjacency([np.asarray(m[lo:i + 1]) for m in data_modalities], modality_types,
                                       k_basis=k_basis, device=device)
            if approach == "sSVDMC":                          # main.py:77-79: randomized TruncatedSVD of this window alone
                red = mo.perform_svd_reduction(fused, reduced_dim, seed, device=device)
            else:
                if swfd is None:                                                                  # main.py:60-62
                    max_norm = float(fused.sum(dim=1, dtype=torch.int64).max().item())            # ||0/1 row||^2 = degree
                    swfd = SeqBasedSWFD(N=window_size, R=max_norm, d=fused.shape[1], sketch_dim=reduced_dim,
                                        schedule=schedule,
                                        device=int(str(device).split(":")[-1]) if ":" in str(device) else 0)
                swfd.fit(fused)                                   # main.py:65-67, the W uint8 rows in one call, device-resident
                red = swfd.get()[0]                               # main.py:70
                red = np.asarray(red.cpu() if hasattr(red, "cpu") else red)
                if red.shape[0] != window_size:                   # main.py:73-76
                    red = red.T
            reduced.append(red)
            clusters = mo.perform_clustering(red, n_clusters, seed, device=device, engine=kmeans)   # main.py:97
            clusters = mo.match_clusters(prev_clusters, clusters, method="hungarian", min_overlap=3,
                                         device=device)                                       # main.py:110
            if clusters is None or len(clusters) == 0:
                clusters = np.full(window_size, 0)
            prev_clusters = clusters
            all_clusters.extend(clusters)
    if swfd is not None:
        swfd.close()
    return np.asarray(all_clusters), np.asarray(all_true), reduced

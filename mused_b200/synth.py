"""Seeded synthetic streams of the shapes BASELINE.json names.

G1 ``lowrank_stream``: the swfd experiment model ``A = S.D.U + Z/zeta``
(/root/reference/swfd/generate_synthetic.py:18-29: m=10, D=diag(1-(i-1)/m),
U = Q^T of a d x m Gaussian, zeta=10) but SEEDED and produced in chunks.
G2 ``event_stream``: bursty multimodal event rows (unit-norm per-modality
blocks around event centres plus noise items), labels returned for NMI/F1.

Pure NumPy on purpose: the CPU baseline and the GPU arm must see the same
bytes.  Values are generated as float32 so that they are exactly
representable on both sides (the reference computes in float64).
"""
from __future__ import annotations

import numpy as np


def lowrank_basis(d, m=10, seed=0):
    rng = np.random.default_rng([seed, 7])
    q, _ = np.linalg.qr(rng.standard_normal((d, m)))
    return np.ascontiguousarray(q.T)                       # m x d


def lowrank_chunks(n, d, m=10, zeta=10.0, seed=0, chunk=65536, scale=1.0):
    """Yield float32 chunks (<= chunk rows) of the n x d stream."""
    U = lowrank_basis(d, m, seed).astype(np.float32)
    D = np.array([1.0 - (i - 1) / m for i in range(m)], dtype=np.float32)
    done = 0
    k = 0
    while done < n:
        r = min(chunk, n - done)
        rng = np.random.default_rng([seed, 11, k])
        S = rng.standard_normal((r, m), dtype=np.float32) * D
        Z = rng.standard_normal((r, d), dtype=np.float32)
        A = S @ U
        A += Z * np.float32(1.0 / zeta)
        if scale != 1.0:
            A *= np.float32(scale)
        yield A
        done += r
        k += 1


def lowrank_stream(n, d, **kw):
    return np.concatenate(list(lowrank_chunks(n, d, **kw)), axis=0)


def row_norm2_max(A):
    A64 = np.asarray(A, dtype=np.float64)
    return float(np.max(np.einsum("ij,ij->i", A64, A64)))


def event_stream(n, blocks=(256, 192, 32, 32), n_events=150, noise_rate=0.95,
                 sigma=0.1, burst=200, seed=0):
    """n x sum(blocks) float32 rows + int labels (0 = noise, 1..E = event id).

    Each event owns one unit-norm centre per modality block; an event item is
    centre + N(0, sigma^2) renormalised per block.  Noise items are uniform
    unit blocks.  Events are active in contiguous bursts of ``burst`` rows.
    """
    rng = np.random.default_rng([seed, 23])
    d = int(sum(blocks))
    centres = []
    for b in blocks:
        c = rng.standard_normal((n_events, b)).astype(np.float32)
        c /= np.linalg.norm(c, axis=1, keepdims=True)
        centres.append(c)
    X = np.empty((n, d), dtype=np.float32)
    labels = np.zeros(n, dtype=np.int64)
    pos = 0
    while pos < n:
        r = min(burst, n - pos)
        active = rng.choice(n_events, size=max(1, min(n_events, 4)), replace=False)
        is_noise = rng.random(r) < noise_rate
        ev = active[rng.integers(0, len(active), size=r)]
        labels[pos:pos + r] = np.where(is_noise, 0, ev + 1)
        off = 0
        for bi, b in enumerate(blocks):
            blk = rng.standard_normal((r, b)).astype(np.float32)
            sig = centres[bi][ev] + np.float32(sigma) * blk
            blk = np.where(is_noise[:, None], blk, sig)
            blk /= np.linalg.norm(blk, axis=1, keepdims=True)
            X[pos:pos + r, off:off + b] = blk
            off += b
        pos += r
    return X, labels, np.concatenate(centres, axis=1)


def multimodal_event_stream(n, n_events=150, noise_rate=0.95, vis_dim=512, txt_dim=768, burst=400, seed=0):
    """MediaEval-SED-shaped synthetic items (BASELINE configs[2]): per item a (latitude, longitude)
    pair, (datetaken, dateupload) epoch seconds, a visual and a text embedding; events own a place,
    a time and a centre per embedding, noise items (label 0, main.py:306) are spread uniformly.
    Returns ([location, time, visual, text], modality_types, labels)."""
    rng = np.random.default_rng([seed, 31])
    ev_loc = np.column_stack([rng.uniform(-50, 50, n_events), rng.uniform(-150, 150, n_events)])
    ev_time = 1.30e9 + rng.uniform(0, 3e7, n_events)
    ev_vis = rng.standard_normal((n_events, vis_dim)).astype(np.float32)
    ev_txt = rng.standard_normal((n_events, txt_dim)).astype(np.float32)
    labels = np.zeros(n, dtype=np.int64)
    is_noise = rng.random(n) < noise_rate
    active = (np.arange(n) // burst) % n_events
    other = rng.integers(0, n_events, n)
    ev = np.where(rng.random(n) < 0.8, active, other)
    labels[~is_noise] = ev[~is_noise] + 1
    loc = np.column_stack([rng.uniform(-60, 60, n), rng.uniform(-180, 180, n)])
    loc[~is_noise] = ev_loc[ev[~is_noise]] + rng.normal(0, 0.05, (int((~is_noise).sum()), 2))
    t0 = 1.30e9 + rng.uniform(0, 3e7, n)
    t0[~is_noise] = ev_time[ev[~is_noise]] + rng.normal(0, 3600, int((~is_noise).sum()))
    tim = np.column_stack([t0, t0 + np.abs(rng.normal(0, 86400, n))])
    vis = rng.standard_normal((n, vis_dim)).astype(np.float32)
    vis[~is_noise] = ev_vis[ev[~is_noise]] + 0.3 * vis[~is_noise]
    txt = rng.standard_normal((n, txt_dim)).astype(np.float32)
    txt[~is_noise] = ev_txt[ev[~is_noise]] + 0.3 * txt[~is_noise]
    return [loc, tim, vis.astype(np.float64), txt.astype(np.float64)], ["location", "time", "visual", "cosine"], labels

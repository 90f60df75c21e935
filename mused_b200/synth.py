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

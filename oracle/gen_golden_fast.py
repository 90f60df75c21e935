"""Fixture for the reference's Fast sketch family (swfd/fastfdwithdump.py:7-138,
swfd/sliding_windows_frequent_directions.py:200-352, swfd/seq_based_swfd.py:209-311) from the UNMODIFIED
reference: FastTimeBasedSWFD on a seeded, time-stamped stream; B^T B of every query, get_size, per-level queue
lengths and (s, t) stamps, plus a FastFrequentDirectionsWithDump run on its own.  Build container only:
    python -m oracle.gen_golden_fast
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import importlib
import os
import sys

import numpy as np
import scipy

from . import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth  # noqa: E402

N, D, ELL, ROWS, QSTEP, R = 200, 48, 8, 700, 100, 4.0


def stream():
    rng = np.random.default_rng(21)
    A = synth.lowrank_stream(ROWS, D, seed=22).astype(np.float64)
    A /= np.linalg.norm(A, axis=1, keepdims=True)
    A *= np.sqrt(rng.uniform(0.3, R, size=(ROWS, 1)))
    A[rng.integers(0, ROWS, 12)] *= 3.0                      # a few heavy rows: raw snapshots at the low levels
    t = 1 + np.cumsum(rng.integers(0, 3, size=ROWS))          # repeated and skipped arrival times
    return A, t.astype(np.int64)


def main():
    swfd_pkg, _, _ = ref_shim.load()
    seq = importlib.import_module("swfd.seq_based_swfd")
    fast = importlib.import_module("swfd.fastfdwithdump")
    A, T = stream()
    sk = seq.FastTimeBasedSWFD(N, R * 9.0, D, ELL)
    grams, sizes, qlens, qstamps = [], [], [], []
    for i in range(ROWS):
        sk.fit(A[i:i + 1], int(T[i]))
        if (i + 1) % QSTEP == 0:
            B = np.asarray(sk.get()[0])
            grams.append(B.T @ B)
            sizes.append(sk.get_size())
            qlens.append([len(sk.levels[j].fd.queue) for j in sk.levels] + [len(sk.levels[j].fd_aux.queue) for j in sk.levels])
            qstamps.append(np.array([[int(e.s), int(e.t)] for j in sk.levels for e in sk.levels[j].fd.queue] or np.zeros((0, 2)), dtype=np.int64))
    # the FD class alone: dumped rows (as B^T B of each dump) and the final stack
    fd = fast.FastFrequentDirectionsWithDump(D, ELL, 6.0)
    dump_at, dump_gram = [], []
    for i in range(300):
        r = fd.fit(A[i:i + 1])
        if r is not None:
            dump_at.append(i)
            dump_gram.append(np.asarray(r).T @ np.asarray(r))
    S = np.asarray(fd.get_sketch())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fast_swfd.npz"), A=A, T=T, N=N, ell=ELL, qstep=QSTEP, R=R * 9.0,
                        n_levels=sk.log_eNR, gram=np.stack(grams), size=np.array(sizes), qlen=np.array(qlens),
                        stamps=np.concatenate(qstamps), stamps_split=np.array([len(q) for q in qstamps]),
                        fd_error=6.0, fd_dump_at=np.array(dump_at), fd_dump_gram=np.stack(dump_gram), fd_final_gram=S.T @ S,
                        fd_final_rows=S.shape[0], fd_energy=fd.max_energy,
                        versions="numpy %s scipy %s" % (np.__version__, scipy.__version__))
    print("levels", sk.log_eNR, "sizes", sizes, "dumps", len(dump_at), "final rows", S.shape[0])


if __name__ == "__main__":
    main()

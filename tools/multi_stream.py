"""Aggregate rows/s of K independent streams batched on one GPU through the native many-streams engine
(mused_swfd_cfg.n_streams; MultiStreamSWFD) at BASELINE configs[1]: d=1024, l=128, N=10000.
Usage: python tools/multi_stream.py [K ...]   (default 1 4 16)"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth                                   # noqa: E402
from mused_b200.swfd import MultiStreamSWFD                    # noqa: E402

D, ELL, N, R = 1024, 128, 10000, 64.0
WINDOWS = 3


def run(K, with_get=True):
    base = [torch.as_tensor(synth.lowrank_stream((WINDOWS + 1) * N, D, seed=k % 4)).cuda() for k in range(min(K, 4))]
    A = torch.stack([base[k % len(base)] for k in range(K)])   # K x rows x d (streams repeat 4 seeds: same cost)
    sk = MultiStreamSWFD(K, N, R, D, ELL)
    sk.fit(A[:, :N])
    for k in range(K):
        sk.get(k)
    torch.cuda.synchronize()
    sk.kernel_ms(enable=1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    tg = 0.0
    for w in range(1, WINDOWS + 1):
        sk.fit(A[:, w * N:(w + 1) * N])
        if with_get:
            torch.cuda.synchronize()
            ta = time.perf_counter()
            for k in range(K):
                sk.get(k)
            tg += time.perf_counter() - ta
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    rows = K * WINDOWS * N
    kms, chunks, _ = sk.kernel_ms(enable=0)
    print("K=%3d  %8.0f rows/s aggregate  (%.1f ms per window of %d x %d rows; get() %.1f ms per window)  per chunk us: %s"
          % (K, rows / ms * 1e3, ms / WINDOWS, K, N, 1e3 * tg / WINDOWS,
             {k: round(1e3 * v / max(chunks, 1), 1) for k, v in kms.items()}), flush=True)
    del sk, A
    torch.cuda.empty_cache()


if __name__ == "__main__":
    ks = [int(x) for x in sys.argv[1:]] or [1, 4, 16]
    for K in ks:
        run(K)

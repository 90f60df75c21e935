"""Isolated time of K1, the TMA-staged append: one fit() call stages up to 8192 rows in one launch (the rest of the
call -- K5, K2, K3, K4 per chunk -- is timed separately by the library's per-kernel events).
Usage: python tools/append_time.py [d ...]      MUSED_STAGE_SIMT=1 -> the SIMT append for comparison."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth                                   # noqa: E402
from mused_b200.swfd import SeqBasedSWFD                       # noqa: E402

peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
for d in [int(a) for a in sys.argv[1:]] or [1024, 4096]:
    ell = 128
    rows = 8192 if d <= 1024 else 8192 * 1024 // d
    A = torch.as_tensor(synth.lowrank_stream(rows * 6, d, seed=1)).cuda()
    sk = SeqBasedSWFD(10000, 64.0 * d / 1024, d, ell, schedule="batched")
    sk.fit(A[:rows])
    torch.cuda.synchronize()
    sk.kernel_ms(enable=1)
    for i in range(1, 6):
        sk.fit(A[i * rows:(i + 1) * rows])
    torch.cuda.synchronize()
    ms, chunks, nrows = sk.kernel_ms(enable=-1)
    k1_us = ms["K1_stage"] / 5 * 1e3
    alg, moved = rows * d * 8.0, rows * d * 12.0
    print("d=%d: %d rows per launch, K1 %.2f us: %.0f GB/s on the algorithmic 8 B/element (%.1f %% of measured HBM peak), "
          "%.0f GB/s on the 12 B/element it moves (%.1f %%)  [%s]"
          % (d, rows, k1_us, alg / k1_us / 1e3, 100 * alg / k1_us / 1e3 / peak, moved / k1_us / 1e3,
             100 * moved / k1_us / 1e3 / peak, "SIMT" if os.environ.get("MUSED_STAGE_SIMT") else "TMA-staged"))
    sk.close()

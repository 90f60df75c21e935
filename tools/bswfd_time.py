"""Per-kernel-class CUDA-event time of the batched engine on a BASELINE config.
Usage: python tools/bswfd_time.py [ell d N rows R]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth                      # noqa: E402
from mused_b200.swfd import SeqBasedSWFD           # noqa: E402

ell, d, N, rows = 128, 1024, 10000, 40000
R = 64.0
if len(sys.argv) >= 5:
    ell, d, N, rows = (int(x) for x in sys.argv[1:5])
if len(sys.argv) >= 6:
    R = float(sys.argv[5])
A = synth.lowrank_stream(rows, d, seed=0)
assert synth.row_norm2_max(A) <= R
Ad = torch.as_tensor(A).cuda()
sk = SeqBasedSWFD(N, R, d, ell, schedule="batched")
sk.fit(Ad[:N])
sk.get()
torch.cuda.synchronize()
for timed in (0, 1):
    if timed:
        sk.kernel_ms(enable=1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(N, rows, N):
        sk.fit(Ad[i:i + N])
        sk.get()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    n = rows - N
    print("%s: %d rows in %.1f ms -> %.0f rows/s (wall %.3fs)" %
          ("with per-kernel events" if timed else "plain", n, ms, n / ms * 1e3, time.perf_counter() - t0))
    if timed:
        k, chunks, r = sk.kernel_ms(enable=0)
        print("chunks %d rows %d" % (chunks, r))
        for name, v in k.items():
            print("  %-10s %9.2f ms total  %8.1f us/chunk" % (name, v, v / max(chunks, 1) * 1e3))
st = sk.debug_state()
print("sweeps(last)", [int(o["energy"]) for o in st], "qlen", [len(o["q_t"]) for o in st])

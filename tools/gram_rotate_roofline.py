"""Tensor-pipe rate of the Gram (K2) and rotate (K4) kernels at growing d (l = 128, 12 FD instances):
CUDA-event time per chunk from the library's own per-kernel timers, dense and issued (3xTF32) TFLOP/s
against the measured bf16 peak / 2 (kind::tf32 runs at half the bf16 rate)."""
import sys, json, ctypes as C
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth

peak_tf32 = json.load(open("MEASURED_PEAKS.json"))["bf16_tflops"] / 2.0
ell, N = 128, 10000
for d in [int(a) for a in sys.argv[1:]] or [1024, 4096, 16384]:
    n = 20 * ell
    A = synth.lowrank_stream(2 * n, d, seed=0, scale=2.0)
    sk = SeqBasedSWFD(N=N, R=64.0, d=d, sketch_dim=ell, schedule="batched")
    X = torch.as_tensor(A).cuda()
    sk.fit(X[:n])
    sk.kernel_ms(enable=1)
    sk.fit(X[n:])
    torch.cuda.synchronize()
    km, chunks, rows = sk.kernel_ms(enable=0)
    S, n2 = 2 * len(sk.thresholds), 2 * ell
    for name, key, flops in (("K2 gram", "K2_gram", S * 2.0 * n2 * n2 * d), ("K4 rotate", "K4_rotate", S * 2.0 * ell * n2 * d)):
        us = km[key] / chunks * 1e3
        print("d=%5d %-9s %8.1f us/chunk  dense %6.1f TFLOP/s  issued(3xTF32) %6.1f TFLOP/s = %4.1f%% of TF32 peak (%.0f)" %
              (d, name, us, flops / us / 1e6, 3 * flops / us / 1e6, 100 * 3 * flops / us / 1e6 / peak_tf32, peak_tf32), flush=True)
    print("d=%5d K3 eig %.1f us/chunk, K1 %.1f, K5 %.1f" % (d, km["K3_eig"] / chunks * 1e3, km["K1_stage"] / chunks * 1e3, km["K5_book"] / chunks * 1e3))
    sk.close()

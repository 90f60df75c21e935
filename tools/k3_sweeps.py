"""Sweeps and time of the batched eigensolve (K3) on the configs[1] stream."""
import sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth, _lib
d, ell, N = 1024, 128, 10000
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
A = synth.lowrank_stream(2 * N, d, seed=0, scale=scale)
sk = SeqBasedSWFD(N=N, R=64.0, d=d, sketch_dim=ell, schedule="batched")
X = torch.as_tensor(A).cuda()
sk.fit(X[:N]); sk.get()
lib = _lib.load()
ms = (C.c_double * 5)(); ch = C.c_int64(0); rows = C.c_int64(0)
lib.mused_swfd_kernel_ms(sk._h, 1, ms, C.byref(ch), C.byref(rows))
sweeps = []
for c in range(0, N, 1280):
    sk.fit(X[N + c:N + c + 1280])
    sweeps.append([int(s["energy"]) for s in sk.debug_state()])
lib.mused_swfd_kernel_ms(sk._h, -1, ms, C.byref(ch), C.byref(rows))
print("sweeps per instance (every 10th chunk):", sweeps[:3], "...", sweeps[-1])
import collections
print("flags (sweeps // 1000: 0 fast path, 5 small-column fallback, 7 failed check):", collections.Counter(int(v) // 1000 for row in sweeps for v in row))
print("mean sweeps %.2f  K3 %.1f us/chunk over %d chunks" % (np.mean(sweeps), 1e3 * ms[3] / max(1, ch.value), ch.value))
B = sk.get()[0]
from oracle import swfd_oracle as so
print("cov_err %.8f" % so.cov_error(A[N:2 * N].astype(np.float64), np.asarray(B.cpu() if hasattr(B, "cpu") else B)))

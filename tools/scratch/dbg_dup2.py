import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD
from oracle.swfd_batched_oracle import BatchedSeqSWFD
ell, d, N, n = 32, 256, 500, 1500
A = synth.lowrank_stream(n, d, seed=4, scale=1.5)
A[100:140] = 0.0
A[300:364] = A[299]
A64 = A.astype(np.float64)
R = max(synth.row_norm2_max(A), 2.0)
gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched")
ora = BatchedSeqSWFD(N, R, d, ell)
np.set_printoptions(precision=4, linewidth=200)
for t in range(0, 448, 32):
    gpu.fit(A[t:t + 32]); ora.fit(A64[t:t + 32])
    if t + 32 < 256:
        continue
    for k in (2, 4):
        g, o = gpu.debug_state()[k], ora.debug_state()[k]
        gs, os_ = np.sort(g["s2"])[::-1], np.sort(o["s2"])[::-1]
        print("t=%d level %d: delta gpu %.6f ora %.6f | nflush %d/%d ndump %d/%d | flags %d" % (t + 32, k // 2, g["delta"], o["delta"], g["n_flush"], o["n_flush"], g["n_dump"], o["n_dump"], g["energy"]))
        print("    gpu s2", gs[:5], "..", gs[-4:], "\n    ora s2", os_[:5], "..", os_[-4:])

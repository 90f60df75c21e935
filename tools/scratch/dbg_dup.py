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
A[700:] = A[:n - 700]
A64 = A.astype(np.float64)
R = max(synth.row_norm2_max(A), 2.0)
gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched")
ora = BatchedSeqSWFD(N, R, d, ell)
print("thresholds", ora.thresholds, "R", R, "norm2 of the repeated row", float(A64[299] @ A64[299]))
step = 32
shown = 0
for t in range(0, n, step):
    gpu.fit(A[t:t + step]); ora.fit(A64[t:t + step])
    for k, (g, o) in enumerate(zip(gpu.debug_state(), ora.debug_state())):
        gs, os_ = np.sort(g["s2"])[::-1], np.sort(o["s2"])[::-1]
        bad = (g["n_flush"], g["n_dump"], g["n_raw"], len(g["q_t"])) != (o["n_flush"], o["n_dump"], o["n_raw"], len(o["q_t"])) or abs(gs[0] - os_[0]) > 1e-3 * max(os_[0], 1.0)
        if bad and shown < 6:
            shown += 1
            print("t=%d inst %d level %d side %d: gpu flush/dump/raw/q %s oracle %s flags %s" % (t + step, k, k // 2, k % 2, (g["n_flush"], g["n_dump"], g["n_raw"], len(g["q_t"])),
                  (o["n_flush"], o["n_dump"], o["n_raw"], len(o["q_t"])), g["energy"]))
            print("   gpu s2 top", gs[:6], "\n   ora s2 top", os_[:6], "thr", ora.thresholds[k // 2])
    if shown >= 6:
        break
print("stopped at", t + step, "flags", [int(g["energy"]) for g in gpu.debug_state()])

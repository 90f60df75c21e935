import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD
from oracle.swfd_batched_oracle import BatchedSeqSWFD
rng = np.random.default_rng(11)
ell, d, N, n = 32, 256, 500, 1500
A = (rng.standard_normal((n, 5)) @ rng.standard_normal((5, d))).astype(np.float32)
A64 = A.astype(np.float64)
R = max(synth.row_norm2_max(A), 2.0)
gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched")
ora = BatchedSeqSWFD(N, R, d, ell)
print("thresholds", ora.thresholds, "R", R)
step = 32
for t in range(0, n, step):
    gpu.fit(A[t:t + step]); ora.fit(A64[t:t + step])
    bad = False
    for k, (g, o) in enumerate(zip(gpu.debug_state(), ora.debug_state())):
        if (g["n_flush"], g["n_dump"], g["n_raw"], len(g["q_t"])) != (o["n_flush"], o["n_dump"], o["n_raw"], len(o["q_t"])):
            print("t=%d inst %d level %d: gpu flush/dump/raw/q %s oracle %s sweeps %s" % (t + step, k, k // 2, (g["n_flush"], g["n_dump"], g["n_raw"], len(g["q_t"])),
                  (o["n_flush"], o["n_dump"], o["n_raw"], len(o["q_t"])), g["energy"]))
            print("   gpu s2 top", np.sort(g["s2"])[::-1][:8], "\n   ora s2 top", np.sort(o["s2"])[::-1][:8], "thr", ora.thresholds[k // 2])
            bad = True
    if bad:
        break
print("done at", t + step, "sweeps", [int(g["energy"]) for g in gpu.debug_state()])

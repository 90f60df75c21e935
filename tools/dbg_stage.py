import sys
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth
ell, d, N = 16, 128, 500
A = synth.lowrank_stream(600, d, seed=0)
R = synth.row_norm2_max(A)
sk = SeqBasedSWFD(N, R, d, ell, schedule="batched")
print("thresholds", sk.thresholds, "R", R)
sk.fit(A[:40])
for s in sk.debug_state()[:4]:
    print({k: s[k] for k in ("level", "energy", "n_flush", "n_dump", "n_raw", "q_t")})

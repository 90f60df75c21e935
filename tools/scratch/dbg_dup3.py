import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD
from oracle.swfd_batched_oracle import BatchedSeqSWFD
ell, d, N = 32, 256, 100000
base = synth.lowrank_stream(400, d, seed=4, scale=1.5)

def run(name, A, R=64.0):
    A = np.ascontiguousarray(A, dtype=np.float32); A64 = A.astype(np.float64)
    gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched"); ora = BatchedSeqSWFD(N, R, d, ell)
    out = []
    for t in range(0, A.shape[0], 32):
        gpu.fit(A[t:t + 32]); ora.fit(A64[t:t + 32])
        worst = 0.0
        for g, o in zip(gpu.debug_state(), ora.debug_state()):
            gs, os_ = np.sort(g["s2"])[::-1], np.sort(o["s2"])[::-1]
            if os_[0] > 0:
                worst = max(worst, float(np.max(np.abs(gs - os_)) / os_[0]))
        out.append("%.1e" % worst)
    print("%-46s max |s2 - oracle| / s2_max per chunk: %s  sweeps %s" % (name, " ".join(out), [int(g["energy"]) for g in gpu.debug_state()][:6]), flush=True)
    gpu.close()

x = base[200]
run("plain 160 rows", base[:160])
run("128 plain + [12 copies, 20 plain]", np.vstack([base[:128], np.tile(x, (12, 1)), base[128:148]]))
run("128 plain + [20 plain, 12 copies]", np.vstack([base[:128], base[128:148], np.tile(x, (12, 1))]))
run("128 plain + [2 copies, 30 plain]", np.vstack([base[:128], np.tile(x, (2, 1)), base[128:158]]))
run("128 plain + 32 copies + 32 plain", np.vstack([base[:128], np.tile(x, (32, 1)), base[128:160]]))
run("128 plain + [16 copies of x, 16 copies of y]", np.vstack([base[:128], np.tile(x, (16, 1)), np.tile(base[201], (16, 1))]))
run("128 plain + [12 near-copies (1e-3 noise), 20 plain]", np.vstack([base[:128], np.tile(x, (12, 1)) + 1e-3 * base[300:312], base[128:148]]))

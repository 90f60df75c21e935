import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD
from oracle.swfd_batched_oracle import BatchedSeqSWFD
ell, d, N = 32, 256, 500
base = synth.lowrank_stream(1500, d, seed=4, scale=1.5)
R = max(synth.row_norm2_max(base), 2.0)

def run(name, A):
    A = np.ascontiguousarray(A, dtype=np.float32); A64 = A.astype(np.float64)
    gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched"); ora = BatchedSeqSWFD(N, R, d, ell)
    print(name)
    for t in range(0, A.shape[0], 32):
        gpu.fit(A[t:t + 32]); ora.fit(A64[t:t + 32])
        if t + 32 < A.shape[0] - 96:
            continue
        errs = []
        for k, (g, o) in enumerate(zip(gpu.debug_state(), ora.debug_state())):
            gs, os_ = np.sort(g["s2"])[::-1], np.sort(o["s2"])[::-1]
            errs.append("%.0e(d%d/%d r%d/%d)" % (np.max(np.abs(gs - os_)) / max(os_[0], 1e-30), g["n_dump"], o["n_dump"], g["n_raw"], o["n_raw"]) if k % 2 == 0 else "")
        print("  t=%d  per level: %s" % (t + 32, " ".join(e for e in errs if e)), flush=True)
    gpu.close()

x = base[299]
run("256 plain + 32 copies + 64 plain", np.vstack([base[:256], np.tile(x, (32, 1)), base[256:320]]))
run("256 plain + 32 copies of 0.3 x + 64 plain", np.vstack([base[:256], np.tile(0.3 * x, (32, 1)), base[256:320]]))
run("256 plain + 32 copies of 3 x + 64 plain", np.vstack([base[:256], np.tile(3 * x, (32, 1)), base[256:320]]))

import sys
import numpy as np
sys.path.insert(0, ".")
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD
from oracle.swfd_batched_oracle import BatchedSeqSWFD
ell, d, N = 32, 256, 500
base = synth.lowrank_stream(1500, d, seed=4, scale=1.5)

def run(name, A, R):
    A = np.ascontiguousarray(A, dtype=np.float32); A64 = A.astype(np.float64)
    gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched"); ora = BatchedSeqSWFD(N, R, d, ell)
    out = []
    for t in range(0, A.shape[0], 32):
        gpu.fit(A[t:t + 32]); ora.fit(A64[t:t + 32])
        worst, where = 0.0, -1
        for k, (g, o) in enumerate(zip(gpu.debug_state(), ora.debug_state())):
            gs, os_ = np.sort(g["s2"])[::-1], np.sort(o["s2"])[::-1]
            if os_[0] > 0:
                e = float(np.max(np.abs(gs - os_)) / os_[0])
                if e > worst: worst, where = e, k
        out.append("%.0e@%d" % (worst, where))
    print("%-40s %s" % (name, " ".join(out[6:])), flush=True)
    gpu.close()

A = base[:448].copy(); A[100:140] = 0; A[300:364] = A[299]
R = max(synth.row_norm2_max(base), 2.0)
run("test stream (zeros + 64 copies)", A, R)
B = base[:448].copy(); B[300:364] = B[299]
run("64 copies, no zero rows", B, R)
C = base[:448].copy(); C[100:140] = 0
run("zero rows only", C, R)
D = base[:448].copy(); D[300:364] = D[299] + 1e-3 * base[1000:1064]
run("64 near-copies", D, R)
E = base[:448].copy(); E[320:364] = E[299]
run("44 copies starting at a chunk boundary", E, R)
F = base[:448].copy(); F[352:364] = F[299]
run("12 copies in the chunk 353..384 only", F, R)

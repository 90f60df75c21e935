"""Time SeqBasedSWFD.get() (batched schedule) at configs[1] after some windows; prints the cov-error.
MUSED_EIG_LEGACY=1 -> old two-sided k_jacobi; MUSED_EIG_NOCOV=1 -> Gram side only."""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth
from oracle import swfd_oracle as so

d, ell, N = 1024, 128, 10000
nwin = int(sys.argv[1]) if len(sys.argv) > 1 else 3
A = synth.lowrank_stream(nwin * N, d, seed=0, scale=2.0)
sk = SeqBasedSWFD(N=N, R=64.0, d=d, sketch_dim=ell, schedule="batched")
X = torch.as_tensor(A).cuda()
for w in range(nwin):
    sk.fit(X[w * N:(w + 1) * N])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    B, s2, vt, delta = sk.get()
    e1.record(); torch.cuda.synchronize()
    B = B.cpu().numpy() if isinstance(B, torch.Tensor) else np.asarray(B)
    err = so.cov_error(A[max(0, (w + 1) * N - N):(w + 1) * N].astype(np.float64), B)
    print("window %d: get() %.2f ms  size=%d  cov_err=%.8f  s2[0]=%.6f s2[-1]=%.6e" %
          (w, e0.elapsed_time(e1), sk.get_size(), err, s2[0], s2[-1]), flush=True)

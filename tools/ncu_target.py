"""Small driver for ncu captures: 20 chunks of the configs[1] stream, one get(), one assign."""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth, matrix_operations as mo
d, ell, N = 1024, 128, 10000
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2560
A = synth.lowrank_stream(N + n, d, seed=0, scale=2.0)
sk = SeqBasedSWFD(N=N, R=64.0, d=d, sketch_dim=ell, schedule="batched")
X = torch.as_tensor(A).cuda()
sk.fit(X[:N])            # fills the snapshot queues so that get() sees a full stack
sk.fit(X[N:N + n])
B = sk.get()[0]
C = torch.randn(150, d, device="cuda")
lab = mo.cosine_assign(X[:N], C, 0.1)
torch.cuda.synchronize()
print("ok", lab.shape)

"""Small driver for ncu captures: a filled window, then n rows, one get(), one assign.
Usage: python tools/ncu_target.py [n=256] [ell=128] [d=1024] [N=10000] [R=64] [schedule=batched]"""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth, matrix_operations as mo
a = sys.argv[1:]
n = int(a[0]) if len(a) > 0 else 256
ell = int(a[1]) if len(a) > 1 else 128
d = int(a[2]) if len(a) > 2 else 1024
N = int(a[3]) if len(a) > 3 else 10000
R = float(a[4]) if len(a) > 4 else 64.0
schedule = a[5] if len(a) > 5 else "batched"
A = synth.lowrank_stream(N + n, d, seed=0)
sk = SeqBasedSWFD(N=N, R=R, d=d, sketch_dim=ell, schedule=schedule)
X = torch.as_tensor(A).cuda()
sk.fit(X[:N])            # fills the snapshot queues so that get() sees a full stack
sk.fit(X[N:N + n])
B = sk.get()[0]
C = torch.randn(150, d, device="cuda")
lab = mo.cosine_assign(X[:N], C, 0.1)
torch.cuda.synchronize()
print("ok", lab.shape)

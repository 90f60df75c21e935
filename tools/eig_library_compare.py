"""Context for K3: the library eigensolver (torch.linalg.eigh -> cuSOLVER syevd / batched syevj) on the same kind of
batch the batched FD shrink solves per chunk: S symmetric (2l x 2l) Gram matrices.  Usage:
python tools/eig_library_compare.py   (prints ms per batch, float32 and float64)"""
import sys
import torch

sys.path.insert(0, ".")
torch.manual_seed(0)
for S, n in ((12, 256), (14, 512)):
    R = torch.randn(S, n, 4 * n, device="cuda", dtype=torch.float64)
    R[:, : n // 2] *= torch.linspace(3.0, 0.3, n // 2, device="cuda", dtype=torch.float64)[None, :, None]
    G = R @ R.transpose(1, 2)
    for dt in (torch.float32, torch.float64):
        A = G.to(dt)
        for _ in range(2):
            torch.linalg.eigh(A)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            torch.linalg.eigh(A)
        e1.record()
        torch.cuda.synchronize()
        print("torch.linalg.eigh (cuSOLVER) batch of %d x (%d x %d) %s: %.2f ms per batch" % (S, n, n, str(dt).split(".")[-1], e0.elapsed_time(e1) / 5))

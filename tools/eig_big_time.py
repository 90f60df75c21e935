"""Time mused_sym_eig (whole-GPU block Jacobi) per n; MUSED_EIG_CB sets the block-width floor."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.linalg import sym_eig

for n in [int(a) for a in sys.argv[1:]] or [256, 512, 1024, 1408]:
    rng = np.random.default_rng(0)
    A = rng.standard_normal((n, 12)) @ rng.standard_normal((12, 1024)) * 3 + rng.standard_normal((n, 1024))
    G = torch.as_tensor(A @ A.T).cuda()
    sym_eig(G)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        ev, V, sw = sym_eig(G, return_sweeps=True)
    e1.record(); torch.cuda.synchronize()
    ref = np.linalg.eigvalsh(G.cpu().numpy())[::-1]
    print("n=%d  %.3f ms/call (incl. scratch alloc)  sweeps=%d  max|dlam|/lmax=%.2e" %
          (n, e0.elapsed_time(e1) / 3, sw, np.max(np.abs(ev.cpu().numpy() - ref)) / ref[0]), flush=True)

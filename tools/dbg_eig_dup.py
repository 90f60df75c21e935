import sys
import numpy as np, torch
sys.path.insert(0, ".")
from oracle import swfd_oracle as so
from mused_b200.linalg import sym_eig
rng = np.random.default_rng(2)
d, ell, N = 40, 6, 50
base = rng.standard_normal((5, d))
A = np.concatenate([np.zeros((7, d)), np.repeat(base, 12, axis=0), np.zeros((3, d)),
                    rng.standard_normal((60, d)), np.repeat(base[:2], 20, axis=0)])
R = float(np.max((A * A).sum(1)))
ora = so.SeqSWFD(N, R, d, ell)
for a in range(0, A.shape[0], 25):
    for r in A[a:a + 25]:
        ora.fit(r[None, :])
    lv = ora.levels[ora.pick_level()]
    q = lv.fd.queue
    if len(q) == 0:
        print("query", a, "no snapshots"); continue
    M = np.vstack([lv.fd.C.get_sketch(), *(e[0] for e in q)])
    for name, G in (("gram", M @ M.T), ("cov", M.T @ M)):
        ev, V, sw = sym_eig(torch.as_tensor(G).cuda(), return_sweeps=True)
        ref = np.linalg.eigvalsh(G)[::-1]
        print("query", a, name, "n=%d sweeps=%d max|dlam|=%.3e lmax=%.3e" % (G.shape[0], sw, np.max(np.abs(ev.cpu().numpy() - ref)), ref[0]),
              "top:", np.round(ev.cpu().numpy()[:4], 4), np.round(ref[:4], 4))

"""First-contact GPU diagnostic (scratch): run the engine on a tiny stream and
print where it departs from the oracle."""
import sys, time, traceback
import numpy as np
sys.path.insert(0, ".")
from oracle import swfd_oracle as so
from mused_b200.swfd import SeqBasedSWFD, FrequentDirectionsWithDump

def diag(cluster, d=24, ell=5, N=60, n=200):
    rng = np.random.default_rng(5)
    A = rng.standard_normal((n, d)) * np.exp(0.7 * rng.standard_normal(n))[:, None]
    R = float(np.max((A * A).sum(1)))
    ora = so.SeqSWFD(N, R, d, ell)
    gpu = SeqBasedSWFD(N, R, d, ell, cluster=cluster)
    bad = 0
    for t in range(n):
        ora.fit(A[t:t+1]); gpu.fit(A[t:t+1])
        if (t + 1) % 10 == 0 or t < 12:
            a, b = ora.debug_state(), gpu.debug_state()
            for x, y in zip(a, b):
                ok = (x["q_s"] == y["q_s"] and x["q_t"] == y["q_t"] and x["pending"] == y["pending"]
                      and abs(x["energy"] - y["energy"]) <= 1e-8 * max(1, abs(x["energy"]))
                      and np.allclose(x["s2"], y["s2"], rtol=1e-8, atol=1e-10))
                if not ok:
                    bad += 1
                    if bad < 6:
                        print("MISMATCH cluster", cluster, "t", t + 1, "level", x["level"])
                        print("  ora", {k: x[k] for k in ("energy", "pending", "last_dump_time")}, x["s2"][:4], x["q_t"][-3:])
                        print("  gpu", {k: y[k] for k in ("energy", "pending", "last_dump_time", "n_flush", "n_dump")}, y["s2"][:4], y["q_t"][-3:])
    a, b = ora.get(), gpu.get()
    print("cluster", cluster, "mismatches", bad, "level", ora.pick_level(), gpu.last_level,
          "sig", np.abs(np.sqrt(a[1]) - np.sqrt(np.maximum(b[1], 0))).max(),
          "BtB", np.abs(a[0].T @ a[0] - b[0].T @ b[0]).max())

for c in (1, 2, 8):
    try:
        diag(c)
    except Exception:
        traceback.print_exc()
try:
    diag(0, d=300, ell=64, N=500, n=1200)
except Exception:
    traceback.print_exc()
# throughput first look
from mused_b200 import synth
import torch
for (d, ell, N, n) in ((512, 64, 2000, 6000), (1024, 128, 10000, 6000)):
    A = synth.lowrank_stream(n, d, seed=0)
    R = synth.row_norm2_max(A)
    for c in (1, 8):
        sk = SeqBasedSWFD(N, R, d, ell, cluster=c)
        Ad = torch.as_tensor(A).cuda()
        sk.fit(Ad[:1000]); torch.cuda.synchronize()
        t0 = time.time(); sk.fit(Ad[1000:]); torch.cuda.synchronize(); t1 = time.time()
        g0 = time.time(); out = sk.get(); g1 = time.time()
        st = sk.debug_state()
        print("d", d, "ell", ell, "cluster", c, "levels", sk.logR, "rows/s %.0f" % ((n - 1000) / (t1 - t0)),
              "get %.1f ms" % ((g1 - g0) * 1e3), "flushes", sum(o["n_flush"] for o in st), "dumps", sum(o["n_dump"] for o in st),
              "level", sk.last_level, "qlen", [len(o["q_s"]) for o in st][:4])

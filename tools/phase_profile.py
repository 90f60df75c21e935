"""Per-FD-instance phase breakdown of the per-row kernel (clock64 counters kept
by the leader thread) on a BASELINE config-2 shaped stream.  GPU box only."""
import sys, time, json
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD

d, ell, N = 1024, 128, 10000
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
cluster = int(sys.argv[2]) if len(sys.argv) > 2 else 0
A = synth.lowrank_stream(n, d, seed=0)
sk = SeqBasedSWFD(N, 64.0, d, ell, cluster=cluster)
Ad = torch.as_tensor(A).cuda()
sk.fit(Ad[:20000]); torch.cuda.synchronize()
p0 = sk.profile()
t0 = time.time(); sk.fit(Ad[20000:]); torch.cuda.synchronize(); t1 = time.time()
p1 = sk.profile()
rows = n - 20000
print("rows/s %.0f  (%.1f us/row)" % (rows / (t1 - t0), 1e6 * (t1 - t0) / rows))
st = sk.debug_state()
names = [k for k in SeqBasedSWFD.PHASES if k != "rows"]
print("slot " + " ".join("%9s" % k for k in names) + "     total  us/row")
for s, (a, b) in enumerate(zip(p0, p1)):
    dlt = [b[k] - a[k] for k in names]
    tot = sum(dlt)
    print("%4d " % s + " ".join("%9.1f" % (v / rows / 1965.0) for v in dlt) + " %9.1f" % (tot / rows / 1965.0))
print("flushes", [o["n_flush"] for o in st])
t0 = time.time(); out = sk.get(); t1 = time.time()
print("get %.1f ms level %d qlen %s" % (1e3 * (t1 - t0), sk.last_level, [len(o["q_s"]) for o in st]))

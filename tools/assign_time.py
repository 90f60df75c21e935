"""Time the fused assignment operator: n x d float32 rows against k centroids (HBM-resident X).
Labels-only cosine calls take the screening path (one TF32 pass + float32 re-scoring of undecided rows);
calls that ask for the score take the float32-exact 3xTF32 kernel.  Two streams: pure noise (the hard case for
screening: small top-2 gaps) and an event stream (rows near a centroid + 30 % noise)."""
import ctypes as C_
import sys, json
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200 import matrix_operations as mo, _lib

d, k = 1024, 150
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
lib = _lib.load()
try:
    import pynvml
    pynvml.nvmlInit()
    _h = pynvml.nvmlDeviceGetHandleByIndex(0)
    sm_mhz = lambda: pynvml.nvmlDeviceGetClockInfo(_h, pynvml.NVML_CLOCK_SM)
except Exception:
    sm_mhz = lambda: -1


def pending():
    v = C_.c_uint64(0)
    _lib.check(lib.mused_assign_pending(C_.byref(v)))
    return v.value


for n in [int(a) for a in sys.argv[1:]] or [10000, 1 << 20]:
    g = torch.Generator(device="cuda").manual_seed(0)
    C = torch.randn(k, d, device="cuda", generator=g)
    for kind in ("noise", "events"):
        X = torch.randn(n, d, device="cuda", generator=g)
        tau = 0.05
        if kind == "events":
            idx = torch.randint(0, k, (n,), device="cuda", generator=g)
            ev = torch.rand(n, device="cuda", generator=g) < 0.7
            X = torch.where(ev[:, None], C[idx] + 0.8 * X, X)
            tau = 0.5
        for score in (False, True):
            for _ in range(3):
                out = mo.cosine_assign(X, C, tau, return_score=score)
            torch.cuda.synchronize()
            p0 = pending()
            reps = 20
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
            evs[0].record()
            for i in range(reps):
                out = mo.cosine_assign(X, C, tau, return_score=score)
                evs[i + 1].record()
            mhz = sm_mhz()                       # sampled while the last launches are still running
            torch.cuda.synchronize()
            lab = out[0] if score else out
            per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(reps))
            ms = evs[0].elapsed_time(evs[reps]) / reps
            spread = "min %.3f / median %.3f / max %.3f ms" % (per[0], per[reps // 2], per[-1])
            frac = (pending() - p0) / reps / n
            gb = (4.0 * n * d + 4.0 * k * d + 4.0 * n) / 1e9
            # check against a float64 torch reference on a slice
            m = min(n, 20000)
            S = torch.nn.functional.normalize(X[:m].double(), dim=1) @ torch.nn.functional.normalize(C.double(), dim=1).T
            top2 = S.topk(2, dim=1).values
            ref = torch.where(top2[:, 0] >= tau, S.argmax(1), torch.full((m,), -1, device="cuda"))
            sure = ((top2[:, 0] - top2[:, 1]) > 2e-5) & ((top2[:, 0] - tau).abs() > 2e-5)
            bad = int((lab[:m].long()[sure] != ref[sure]).sum())
            print("n=%d %-6s %-22s %.3f ms  %.1f GB/s (%.1f%% of measured HBM peak)  re-scored %.2f%%  mismatches(sure rows)=%d/%d  SM %d MHz  [%s]" %
                  (n, kind, "3xTF32 (with score)" if score else "screen + re-score", ms, gb / ms * 1e3,
                   100 * gb / ms * 1e3 / peak, 100 * frac, bad, int(sure.sum()), mhz, spread), flush=True)

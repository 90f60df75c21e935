"""Time the fused assignment kernel: n x d float32 rows against k centroids (HBM-resident X)."""
import sys, json
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200 import matrix_operations as mo

d, k = 1024, 150
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
for n in [int(a) for a in sys.argv[1:]] or [10000, 1 << 20]:
    g = torch.Generator(device="cuda").manual_seed(0)
    X = torch.randn(n, d, device="cuda", generator=g)
    C = torch.randn(k, d, device="cuda", generator=g)
    for _ in range(3):
        lab = mo.cosine_assign(X, C, 0.05)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        lab = mo.cosine_assign(X, C, 0.05)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gb = (4.0 * n * d + 4.0 * k * d + 8.0 * n) / 1e9
    # check against a float64 torch reference on a slice
    m = min(n, 20000)
    S = torch.nn.functional.normalize(X[:m].double(), dim=1) @ torch.nn.functional.normalize(C.double(), dim=1).T
    top2 = S.topk(2, dim=1).values
    ref = torch.where(top2[:, 0] >= 0.05, S.argmax(1), torch.full((m,), -1, device="cuda"))
    sure = ((top2[:, 0] - top2[:, 1]) > 2e-5) & ((top2[:, 0] - 0.05).abs() > 2e-5)
    bad = int((lab[:m].long()[sure] != ref[sure]).sum())
    print("n=%d: %.3f ms  %.1f GB/s (%.1f%% of measured HBM peak)  %.1f TFLOP/s dense  mismatches(sure rows)=%d/%d" %
          (n, ms, gb / ms * 1e3, 100 * gb / ms * 1e3 / peak, 2.0 * n * k * d / ms / 1e9, bad, int(sure.sum())), flush=True)

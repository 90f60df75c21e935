"""Bring-up diagnostics of the batched-schedule engine on a GPU box: runs the CUDA
engine and the float64 batched oracle on the same seeded stream and prints where
they differ (schedule, sigma^2, covariance error).  Not a test; tests/ hold the
assertions.  Usage: python tools/bswfd_check.py [ell d N n scale]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mused_b200 import synth                                   # noqa: E402
from mused_b200.swfd import SeqBasedSWFD                        # noqa: E402
from oracle import swfd_oracle as so                            # noqa: E402
from oracle.swfd_batched_oracle import BatchedSeqSWFD           # noqa: E402


def run(ell, d, N, n, scale, qstep=None, seed=0):
    A = synth.lowrank_stream(n, d, seed=seed, scale=scale)
    A64 = A.astype(np.float64)
    R = synth.row_norm2_max(A)
    qstep = qstep or N // 2
    gpu = SeqBasedSWFD(N, R, d, ell, schedule="batched")
    ora = BatchedSeqSWFD(N, R, d, ell)
    print("== ell=%d d=%d N=%d n=%d R=%.1f levels=%d" % (ell, d, N, n, R, gpu.logR), flush=True)
    t0 = time.time()
    worst_sig = 0.0
    for t in range(0, n, qstep):
        blk = A[t:t + qstep]
        gpu.fit(blk)
        ora.fit(A64[t:t + qstep])
        te = t + blk.shape[0]
        if te < N:
            continue
        sg, s2g, vtg, dg = gpu.get()
        so_, s2o, _, do = ora.get()
        Aw = A64[te - N:te]
        eg, eo = so.cov_error(Aw, sg), so.cov_error(Aw, so_)
        a, b = np.sqrt(np.maximum(s2o, 0)), np.sqrt(np.maximum(s2g, 0))
        big = a >= 0.05 * a.max()
        rel = np.max(np.abs(a[big] - b[big]) / a[big])
        worst_sig = max(worst_sig, rel)
        stg, sto = gpu.debug_state(), ora.debug_state()
        sched_ok = all(g["q_s"] == o["q_s"] and g["q_t"] == o["q_t"] and g["n_flush"] == o["n_flush"]
                       and g["n_dump"] == o["n_dump"] and g["n_raw"] == o["n_raw"]
                       and g["last_dump_time"] == o["last_dump_time"] for g, o in zip(stg, sto))
        print("t=%d level %d/%d cov_err gpu %.6f oracle %.6f (rel diff %.2e)  sigma rel %.2e  abs small %.2e  "
              "schedule %s size %d/%d sweeps %s" %
              (te, gpu.last_level, ora.pick_level(), eg, eo, abs(eg - eo) / eo, rel,
               np.max(np.abs(a[~big] - b[~big]), initial=0.0) / a.max(),
               "same" if sched_ok else "DIFF", gpu.get_size(), ora.get_size(),
               [int(g["energy"]) for g in stg]), flush=True)
        if not sched_ok:
            for i, (g, o) in enumerate(zip(stg, sto)):
                if g["q_t"] != o["q_t"] or g["n_dump"] != o["n_dump"] or g["n_raw"] != o["n_raw"]:
                    print("   inst %d: flush %d/%d dump %d/%d raw %d/%d qlen %d/%d last %d/%d" %
                          (i, g["n_flush"], o["n_flush"], g["n_dump"], o["n_dump"], g["n_raw"], o["n_raw"],
                           len(g["q_t"]), len(o["q_t"]), g["last_dump_time"], o["last_dump_time"]))
                    s2d = np.abs(np.sort(g["s2"])[::-1] - np.sort(o["s2"])[::-1]).max() / max(o["s2"].max(), 1e-300)
                    print("      s2 max rel-to-max diff %.2e" % s2d)
    print("   %.1fs, worst sigma rel %.2e" % (time.time() - t0, worst_sig), flush=True)
    gpu.close()


if __name__ == "__main__":
    if len(sys.argv) >= 5:
        ell, d, N, n = (int(x) for x in sys.argv[1:5])
        run(ell, d, N, n, float(sys.argv[5]) if len(sys.argv) > 5 else 1.0)
    else:
        run(16, 128, 500, 2000, 1.0)
        run(16, 128, 500, 2000, 3.0)
        run(32, 256, 700, 2100, 2.0)
        run(64, 512, 1000, 2500, 2.0)
        run(128, 1024, 2000, 4000, 2.0)

"""2+ ranks, NCCL: column-sharded SWFD with the allreduces inside the C library
(mused_b200.sharded.NcclShardedSWFD) against the single-GPU batched engine on the same stream.
  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/sharded_check.py [ell d N n]
Prints one line per query (rank 0) and exits non-zero on a mismatch (tests/test_sharded_gpu.py runs it)."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth                                   # noqa: E402
from mused_b200.sharded import NcclShardedSWFD                 # noqa: E402
from mused_b200.swfd import SeqBasedSWFD                        # noqa: E402
from oracle import swfd_oracle as so                            # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    ell, d, N, n = (int(x) for x in sys.argv[1:5]) if len(sys.argv) >= 5 else (64, 2048, 2000, 6000)
    A = synth.lowrank_stream(n, d, seed=2, scale=2.0)
    R = synth.row_norm2_max(A)
    sh = NcclShardedSWFD(N, R, d, ell, device=lr)
    one = SeqBasedSWFD(N, R, d, ell, device=lr, schedule="batched") if rank == 0 else None
    Ad = torch.as_tensor(A).cuda()
    ok = True
    t_sh = 0.0
    for t in range(0, n, N // 2):
        blk = Ad[t:t + N // 2]
        torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
        sh.fit(blk)
        torch.cuda.synchronize(); t_sh += time.perf_counter() - t0
        if one is not None:
            one.fit(blk)
        te = t + blk.shape[0]
        if te < N:
            continue
        sk, s2, vt, delta = sh.get()
        if rank == 0:
            sk1, s21, _, _ = one.get()
            Aw = A[te - N:te].astype(np.float64)
            e_sh, e_1 = so.cov_error(Aw, sk), so.cov_error(Aw, sk1)
            a, b = np.sqrt(np.maximum(s21, 0)), np.sqrt(np.maximum(s2, 0))
            big = a >= 0.05 * a.max()
            rel = float(np.max(np.abs(a[big] - b[big]) / a[big]))
            same = all(x["q_t"] == y["q_t"] and x["q_s"] == y["q_s"] for x, y in zip(sh.debug_state(), one.debug_state()))
            good = same and rel < 5e-6 and abs(e_sh - e_1) < 5e-5 * e_1 and sh.last_level == one.last_level
            ok = ok and good
            print("t=%d world=%d cov_err sharded %.6f single %.6f sigma rel %.2e level %d/%d schedule %s -> %s"
                  % (te, world, e_sh, e_1, rel, sh.last_level, one.last_level, "same" if same else "DIFF",
                     "ok" if good else "MISMATCH"), flush=True)
        else:
            sh.debug_state()
    if rank == 0:
        na, nb = sh.comm_stats()
        print("sharded fit: %d rows in %.3fs -> %.0f rows/s (d=%d over %d GPUs, %d allreduces, %.1f MB)"
              % (n, t_sh, n / t_sh, d, world, na, nb / 1e6), flush=True)
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(int(flag.item() > 0))


if __name__ == "__main__":
    main()

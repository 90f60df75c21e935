"""GPU: the column-sharded driver on the CUDA shard engine.  Single process
(world size 1) checks that the phase C ABI (stage / chunk_gram / chunk_finish /
get_begin / get_finish) reproduces the plain fit()/get() path: identical
schedule and snapshot stamps, and the same query result up to row signs and
round-off (the single-GPU get() eigensolves the smaller of the stack's Gram
and covariance, the sharded one always the allreduced Gram); the
2-rank NCCL run is tools/sharded_check.py under torchrun (needs 2 GPUs)."""
import numpy as np
import pytest

from mused_b200 import synth

pytestmark = pytest.mark.gpu


def test_phase_api_equals_fit_get_single_rank():
    from mused_b200.sharded import ColumnShardedSWFD
    from mused_b200.swfd import SeqBasedSWFD
    ell, d, N, n = 32, 256, 500, 1300
    A = synth.lowrank_stream(n, d, seed=6, scale=2.0)
    R = synth.row_norm2_max(A)
    a = SeqBasedSWFD(N, R, d, ell, schedule="batched")
    b = ColumnShardedSWFD(N, R, d, ell)
    pos = 0
    for step in (7, 493, 300, 500):
        a.fit(A[pos:pos + step])
        b.fit(A[pos:pos + step])
        pos += step
        if pos >= N:
            ga, gb = a.get(), b.get()
            smax = float(np.max(ga[1]))
            np.testing.assert_allclose(ga[1], gb[1], rtol=0, atol=1e-10 * smax)
            for k in (0, 2):
                sign = np.sign(np.einsum("ij,ij->i", ga[k], gb[k]))
                sign[sign == 0] = 1.0
                scale = np.sqrt(smax) if k == 0 else 1.0
                keep = ga[1] > 1e-6 * smax          # directions of (numerically) zero rows are arbitrary
                np.testing.assert_allclose((ga[k] * sign[:, None])[keep], gb[k][keep], rtol=0, atol=1e-6 * scale)
            np.testing.assert_allclose(ga[0].T @ ga[0], gb[0].T @ gb[0], rtol=0, atol=1e-9 * smax)
            assert ga[3] == gb[3] and a.last_level == b.last_level and a.get_size() == b.get_size()
    for x, y in zip(a.debug_state(), b.debug_state()):
        assert x["q_t"] == y["q_t"] and x["q_s"] == y["q_s"]
    a.close()
    b.close()


def test_nccl_sharded_two_ranks_equals_single_gpu():
    """2 ranks over NCCL (allreduces posted from inside the C ABI): same schedule, sigma within 5e-6, cov-error
    within 5e-5 relative of the single-GPU batched engine.  Needs 2 GPUs (skipped otherwise)."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tools", "sharded_check.py"), "64", "2048", "2000", "5000"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "MISMATCH" not in p.stdout and "-> ok" in p.stdout

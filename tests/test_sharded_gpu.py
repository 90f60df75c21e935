"""GPU: the column-sharded driver on the CUDA shard engine.  Single process
(world size 1) checks that the phase C ABI (stage / chunk_gram / chunk_finish /
get_begin / get_finish) reproduces the plain fit()/get() path bit for bit; the
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
            np.testing.assert_array_equal(ga[0], gb[0])
            np.testing.assert_array_equal(ga[1], gb[1])
            np.testing.assert_array_equal(ga[2], gb[2])
            assert ga[3] == gb[3] and a.last_level == b.last_level and a.get_size() == b.get_size()
    for x, y in zip(a.debug_state(), b.debug_state()):
        assert x["q_t"] == y["q_t"] and x["q_s"] == y["q_s"]
    a.close()
    b.close()

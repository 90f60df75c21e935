"""CPU: the batched-schedule oracle (oracle/swfd_batched_oracle.py) against the
reference-schedule oracle (bit-identical port of the reference) on the same seeded
streams: covariance error comparable and under the bound, same level choice, and
independence of how the caller slices the rows over fit() calls."""
import numpy as np

from mused_b200 import synth
from oracle import swfd_oracle as so
from oracle.swfd_batched_oracle import BatchedSeqSWFD


def _run(d, ell, N, n, scale, q):
    A = synth.lowrank_stream(n, d, seed=1, scale=scale).astype(np.float64)
    R = synth.row_norm2_max(A)
    ref, bat = so.SeqSWFD(N, R, d, ell), BatchedSeqSWFD(N, R, d, ell)
    out = []
    for t in range(n):
        ref.fit(A[t:t + 1])
        bat.fit(A[t:t + 1])
        if (t + 1) % q == 0 and t + 1 >= N:
            Aw = A[t + 1 - N:t + 1]
            out.append((so.cov_error(Aw, ref.get()[0]), so.cov_error(Aw, bat.get()[0]),
                        ref.pick_level(), bat.pick_level()))
    return out


def test_batched_schedule_cov_error_tracks_reference_schedule():
    for (d, ell, N, n, scale) in [(64, 8, 200, 800, 1.0), (96, 16, 300, 900, 3.0)]:
        for er, eb, lr, lb in _run(d, ell, N, n, scale, N // 2):
            assert er <= 1.0 / ell and eb <= 1.0 / ell          # swfd/test_synthetic.py:84-88 bound
            assert eb <= 1.25 * er, (er, eb)
            assert lr == lb


def test_batched_oracle_independent_of_slicing_and_get_closes_chunk():
    rng = np.random.default_rng(2)
    d, ell, N, n = 48, 8, 120, 500
    A = rng.standard_normal((n, d)) * np.exp(0.5 * rng.standard_normal(n))[:, None]
    R = synth.row_norm2_max(A)
    a, b = BatchedSeqSWFD(N, R, d, ell), BatchedSeqSWFD(N, R, d, ell)
    for t in range(n):
        a.fit(A[t:t + 1])
    pos = 0
    for step in (3, 50, 7, 240, 200):
        b.fit(A[pos:pos + step])
        pos += step
    assert pos == n
    sa, sb = a.get(), b.get()
    np.testing.assert_array_equal(sa[1], sb[1])
    np.testing.assert_array_equal(sa[0], sb[0])
    assert a.get_size() == b.get_size()
    assert a.done == a.time == n and not a.pend              # get() folded the open chunk
    for x, y in zip(a.debug_state(), b.debug_state()):
        assert x["q_t"] == y["q_t"] and x["q_s"] == y["q_s"]
        # stamps: s <= t for every snapshot, expiry leaves nothing with t + N <= time
        assert all(s <= t for s, t in zip(x["q_s"], x["q_t"]))
        assert all(t + N > n for t in x["q_t"])


def test_batched_oracle_zero_levels_raises_like_reference():
    import pytest
    sk = BatchedSeqSWFD(100, 1.0, 16, 4)
    sk.fit(np.ones((3, 16)) * 0.1)
    with pytest.raises(KeyError):
        sk.get()

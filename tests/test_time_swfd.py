"""TimeBasedSWFD (swfd/seq_based_swfd.py:106-207): explicit arrival times, repeated stamps, gaps of
several windows.  Golden outputs of the reference: tests/golden/time_swfd.npz (oracle/gen_time_golden.py)."""
import os

import numpy as np
import pytest

from oracle import swfd_oracle as so

GOLD = os.path.join(os.path.dirname(__file__), "golden", "time_swfd.npz")


def _run(sk, g, block):
    A, times = g["A"], g["times"]
    out, nxt = [], 0
    q_rows = [int(v) for v in g["q_rows"]]
    pos = 0
    for qr in q_rows:
        while pos < qr:
            m = min(block, qr - pos)
            if block == 1:
                sk.fit(A[pos:pos + 1], times[pos])
            else:
                sk.fit(A[pos:pos + m], times[pos:pos + m])
            pos += m
        size = sk.get_size()
        sketch, s2, _, _ = sk.get()
        out.append((np.asarray(sketch), np.asarray(s2), size))
    return out


def test_time_oracle_matches_reference_golden():
    g = np.load(GOLD, allow_pickle=False)
    N, ell = (int(v) for v in g["params"])
    ora = so.TimeSWFD(N, float(g["R"]), g["A"].shape[1], ell)
    for k, (sk, s2, size) in enumerate(_run(ora, g, 1)):
        assert np.array_equal(sk, g["sketch"][k]) and np.array_equal(s2, g["s2"][k]) and size == int(g["size"][k])


@pytest.mark.gpu
@pytest.mark.parametrize("block", [1, 37])
def test_time_swfd_device_matches_reference_golden(block):
    from mused_b200.swfd import TimeBasedSWFD
    g = np.load(GOLD, allow_pickle=False)
    N, ell = (int(v) for v in g["params"])
    dev = TimeBasedSWFD(N, float(g["R"]), g["A"].shape[1], ell)
    assert len(dev.thresholds) == int(np.ceil(np.log2(N * float(g["R"]) / ell)))
    for k, (sk, s2, size) in enumerate(_run(dev, g, block)):
        ref_sk, ref_s2 = g["sketch"][k], g["s2"][k]
        assert size == int(g["size"][k])
        a, b = np.sqrt(np.maximum(s2, 0)), np.sqrt(np.maximum(ref_s2, 0))
        big = b >= 1e-3 * b.max()
        assert np.max(np.abs(a[big] - b[big]) / b[big]) < 1e-8
        sc = np.abs(ref_sk.T @ ref_sk).max()
        assert np.abs(sk.T @ sk - ref_sk.T @ ref_sk).max() <= 1e-8 * sc

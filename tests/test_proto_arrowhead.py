"""CPU: numerics of the device algorithm (tests/proto_arrowhead.py) -- the
arrowhead/secular eigen-update that replaces the reference's thin SVD --
against LAPACK and against the oracle's flush/dump schedule."""
import numpy as np

import proto_arrowhead as pa
from oracle import swfd_oracle as so


def _check(d, w, alpha):
    n = len(d)
    M = np.zeros((n + 1, n + 1))
    M[:n, :n] = np.diag(d)
    M[:n, n] = w
    M[n, :n] = w
    M[n, n] = alpha
    lam, U = pa.arrowhead_eig(d, w, alpha)
    ref = np.linalg.eigvalsh(M)[::-1]
    sc = max(1e-300, np.abs(ref).max())
    assert np.abs(np.sort(lam)[::-1] - ref).max() / sc < 1e-13
    assert np.abs(U.T @ U - np.eye(n + 1)).max() < 1e-13
    assert np.abs(U @ np.diag(lam) @ U.T - M).max() / sc < 1e-13


def test_arrowhead_eig_random_and_degenerate():
    rng = np.random.default_rng(0)
    for trial in range(120):
        n = int(rng.integers(1, 40))
        kind = trial % 6
        d = np.sort(rng.random(n))[::-1] * 10
        w = rng.standard_normal(n)
        if kind == 1:
            d[n // 2:] = 0
            w[n // 2:] = 0
        if kind == 2:
            d[:] = np.round(d)
        if kind == 3:
            w *= 1e-9
        if kind == 4:
            d[1:] = d[0] * (1 + 1e-15 * rng.standard_normal(n - 1))
        if kind == 5:
            d = rng.permutation(d)
            w[rng.integers(0, n)] = 1e-20
        _check(d, w, abs(rng.standard_normal()) * 5)


def test_arrowfd_follows_reference_schedule():
    rng = np.random.default_rng(5)
    A = rng.standard_normal((260, 24)) * np.exp(0.7 * rng.standard_normal(260))[:, None]
    R = float(np.max((A * A).sum(1)))

    def run(cls):
        keep = so.FDDump
        so.FDDump = cls
        try:
            sk = so.SeqSWFD(60, R, 24, 5)
            out = []
            for t in range(A.shape[0]):
                sk.fit(A[t:t + 1])
                if (t + 1) % 65 == 0:
                    st = sk.debug_state()
                    out.append((st, sk.pick_level(), sk.get()))
            return out
        finally:
            so.FDDump = keep

    for (sa, ja, ga), (sb, jb, gb) in zip(run(so.FDDump), run(pa.ArrowFD)):
        assert ja == jb
        for a, b in zip(sa, sb):
            assert a["q_s"] == b["q_s"] and a["q_t"] == b["q_t"] and a["pending"] == b["pending"]
        s_a, s_b = np.sqrt(ga[1]), np.sqrt(np.maximum(gb[1], 0))
        big = s_a >= 1e-3 * s_a.max()
        assert np.max(np.abs(s_a[big] - s_b[big]) / s_a[big]) < 1e-9

"""tests/golden/time_swfd.npz: the UNMODIFIED reference's TimeBasedSWFD (swfd/seq_based_swfd.py:106-207)
on a seeded stream with repeated arrival times, gaps and one gap of several windows, side by side with
the oracle restatement (asserted bit-identical).  Build container only: python -m oracle.gen_time_golden
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import os

import numpy as np
import scipy

from . import ref_shim, swfd_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "time_swfd.npz")


def make(n=900, d=48, seed=0, N=120):
    rng = np.random.default_rng([seed, 5])
    A = rng.standard_normal((n, d)) * np.exp(rng.standard_normal(n))[:, None] * 0.7
    t, times = 0, []
    for i in range(n):
        r = rng.random()
        t += 0 if (r < 0.1 and i > 0) else (1 if r < 0.85 else int(rng.integers(2, 30)))
        if i == 500:
            t += 3 * N + 5
        times.append(max(t, 1))
    return A, np.asarray(times, dtype=np.uint64)


def main():
    swfd_pkg, mo, me = ref_shim.load()
    from swfd.seq_based_swfd import TimeBasedSWFD
    N, ell = 120, 8
    A, times = make(N=N)
    R = float(np.max((A * A).sum(1)))
    ref, ora = TimeBasedSWFD(N, R, A.shape[1], ell), so.TimeSWFD(N, R, A.shape[1], ell)
    rec = []
    for i in range(A.shape[0]):
        ref.fit(A[i:i + 1], times[i])
        ora.fit(A[i:i + 1], times[i])
        if (i + 1) % 90 == 0:
            a, b = ref.get(), ora.get()
            assert all(np.array_equal(np.asarray(x), np.asarray(y)) for x, y in zip(a, b))
            assert ref.get_size() == ora.get_size()
            rec.append((i + 1, np.array(a[0]), np.array(a[1]), ref.get_size()))
    np.savez_compressed(OUT, A=A, times=times, params=np.array([N, ell]), R=R,
                        q_rows=np.array([r[0] for r in rec]), sketch=np.stack([r[1] for r in rec]),
                        s2=np.stack([r[2] for r in rec]), size=np.array([r[3] for r in rec]),
                        versions="numpy %s scipy %s" % (np.__version__, scipy.__version__))
    print("wrote", OUT, len(rec), "queries, levels", ref.log_eNR)


if __name__ == "__main__":
    main()

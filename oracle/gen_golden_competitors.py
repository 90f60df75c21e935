"""Fixtures for the competitor sketches LM-FD and DI-FD (swfd/lmfd.py:18-117, swfd/difd.py:41-142) from the
UNMODIFIED reference: B^T B of every query, get_size, on a seeded stream.  Build container only:
    python -m oracle.gen_golden_competitors
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import importlib
import os
import sys

import numpy as np
import scipy

from . import ref_shim, swfd_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth  # noqa: E402

N, D, ELL, ROWS, QSTEP, R = 300, 64, 8, 900, 150, 2.0


def stream():
    A = synth.lowrank_stream(ROWS, D, seed=11).astype(np.float64)
    rng = np.random.default_rng(12)
    A /= np.linalg.norm(A, axis=1, keepdims=True)
    A *= np.sqrt(rng.uniform(0.5, R, size=(ROWS, 1)))          # 0.5 <= ||x||^2 <= R
    return A


def main():
    ref_shim.load()
    lm = importlib.import_module("swfd.lmfd").LMFD(N, D, ELL)
    di = importlib.import_module("swfd.difd").DIFD(N, D, ELL, R)
    A = stream()
    out = dict(A=A, N=N, ell=ELL, qstep=QSTEP, R=R, versions="numpy %s scipy %s" % (np.__version__, scipy.__version__))
    for name, sk in (("lmfd", lm), ("difd", di)):
        grams, sizes, errs = [], [], []
        for t in range(ROWS):
            sk.fit(A[t:t + 1])
            if (t + 1) % QSTEP == 0:
                B = np.asarray(sk.get()[0])
                grams.append(B.T @ B)
                sizes.append(sk.get_size())
                errs.append(so.cov_error(A[max(0, t + 1 - N):t + 1], B))
        out[name + "_gram"] = np.stack(grams)
        out[name + "_size"] = np.array(sizes)
        out[name + "_cov_err"] = np.array(errs)
        print(name, "sizes", sizes, "cov_err", np.round(errs, 5))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "competitors.npz"), **out)


if __name__ == "__main__":
    main()

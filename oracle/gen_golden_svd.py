"""Reference fixture for perform_svd_reduction (matrix_operations.py:143-147), the reducer of the sSVDMC approaches.

Runs the UNMODIFIED reference function (through oracle/ref_shim.py) on seeded 0/1 adjacency-like matrices and stores
the inputs' generator arguments and the reference outputs.  Build container only (needs /root/reference):
    python -m oracle.gen_golden_svd
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os

import numpy as np
import sklearn

from . import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "svd_reduce.npz")
CASES = [dict(W=300, blocks=6, p_in=0.35, p_out=0.01, seed=1, reduced_dim=20, rs=3),
         dict(W=517, blocks=9, p_in=0.20, p_out=0.02, seed=2, reduced_dim=50, rs=0),
         dict(W=64, blocks=3, p_in=0.5, p_out=0.05, seed=3, reduced_dim=100, rs=7)]     # reduced_dim > W - 1: clipped


def block_adjacency(W, blocks, p_in, p_out, seed):
    """Directed 0/1 matrix with community structure (what the fused kNN adjacency of an event window looks like)."""
    rng = np.random.default_rng([seed, 77])
    sizes = rng.multinomial(W, np.arange(1, blocks + 1) / np.arange(1, blocks + 1).sum())
    lab = np.repeat(np.arange(blocks), sizes)
    P = np.where(lab[:, None] == lab[None, :], p_in, p_out)
    M = (rng.random((W, W)) < P).astype(np.uint8)
    np.fill_diagonal(M, 0)
    return M


def main():
    _, mo, _ = ref_shim.load()
    out = dict(versions="numpy %s sklearn %s" % (np.__version__, sklearn.__version__), n_cases=len(CASES))
    for i, c in enumerate(CASES):
        M = block_adjacency(c["W"], c["blocks"], c["p_in"], c["p_out"], c["seed"])
        X = mo.perform_svd_reduction(M.astype(int), c["reduced_dim"], c["rs"])            # the reference call (main.py:79)
        out["case%d_args" % i] = np.array([c["W"], c["blocks"], c["seed"], c["reduced_dim"], c["rs"]])
        out["case%d_p" % i] = np.array([c["p_in"], c["p_out"]])
        out["case%d_sum" % i] = int(M.sum())
        out["case%d_X" % i] = X
        print("case %d: W=%d -> reduced %s, column norms %s ..." % (i, c["W"], X.shape, np.round(np.linalg.norm(X, axis=0)[:5], 4)))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()

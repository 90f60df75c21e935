"""Reference fixture for the string modalities of create_adjacency_matrix: "username" (matrix_operations.py:55-72)
and "tags" (:74-89).  Runs the UNMODIFIED reference (through oracle/ref_shim.py) on seeded random string columns and
stores the columns with the reference's adjacency matrices.  Build container only (needs /root/reference):
    python -m oracle.gen_golden_strings
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import contextlib
import io
import os

import numpy as np

from . import ref_shim, ops_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "string_modalities.npz")


def columns(W, seed):
    rng = np.random.default_rng([seed, 99])
    pool = [""] * 3 + ["user%02d" % i for i in range(25)]
    users = np.array([pool[rng.integers(0, len(pool))] for _ in range(W)])
    letters = list("abcdefghijklmnopqrstuvwx ,")
    tags = np.array(["".join(rng.choice(letters, size=rng.integers(0, 14))) for _ in range(W)])
    return users, tags


def main():
    _, mo, _ = ref_shim.load()
    out = {}
    for case, (W, k, seed) in enumerate([(300, 10, 1), (97, 50, 2)]):
        users, tags = columns(W, seed)
        with contextlib.redirect_stdout(io.StringIO()):
            au = mo.create_adjacency_matrix(users.astype(object).reshape(-1, 1), "username", k_basis=k)
            at = mo.create_adjacency_matrix(tags.astype(object).reshape(-1, 1), "tags", k_basis=k)
        ora_u = ops_oracle.username_adjacency(users)
        ora_t, det = ops_oracle.tags_adjacency(tags, k)
        assert np.array_equal(au, ora_u)
        assert np.array_equal(at[det], ora_t[det])
        out.update({"c%d_args" % case: np.array([W, k, seed]), "c%d_users" % case: users, "c%d_tags" % case: tags,
                    "c%d_adj_users" % case: au.astype(np.uint8), "c%d_adj_tags" % case: at.astype(np.uint8)})
        print("case %d: W=%d k=%d: username edges %d, tags edges %d, rows determined by the reference %d / %d"
              % (case, W, k, int(au.sum()), int(at.sum()), int(det.sum()), W))
    out["n_cases"] = 2
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()

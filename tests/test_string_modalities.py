"""The string modalities of create_adjacency_matrix -- "username" (matrix_operations.py:55-72) and "tags" (:74-89,
Jaccard of the tag sets, top-k_basis) -- against tests/golden/string_modalities.npz, produced by the UNMODIFIED
reference (oracle/gen_golden_strings.py).

"username" is determined completely.  "tags": the reference ranks with np.argsort, whose order inside ties is left
to numpy's introsort; a row is determined by the reference when its k-th and (k+1)-th similarities differ.  Those
rows must be identical; on the others the chosen neighbours must be a valid top-k (same similarity values).
CPU test: the oracle restatement against the fixture.  GPU test: the device operators (C ABI mused_group_adjacency /
mused_jaccard_topk_adjacency) against the fixture and, bit for bit, against the oracle (both break ties by index).
"""
import os

import numpy as np
import pytest

from oracle import ops_oracle


def _cases(golden_dir):
    p = os.path.join(golden_dir, "string_modalities.npz")
    if not os.path.exists(p):
        pytest.skip("fixture missing")
    g = np.load(p)
    for c in range(int(g["n_cases"])):
        W, k, _ = (int(v) for v in g["c%d_args" % c])
        yield W, k, g["c%d_users" % c], g["c%d_tags" % c], g["c%d_adj_users" % c], g["c%d_adj_tags" % c]


def _check_tags(adj, ref, tags, k):
    ora, det = ops_oracle.tags_adjacency(tags, k)
    assert np.array_equal(adj[det], ref[det]), "rows the reference determines differ"
    S, valid = ops_oracle.jaccard_matrix(tags)
    full = np.full((len(tags), len(tags)), -2.0)
    full[np.ix_(valid, valid)] = S
    for i in np.where(~det)[0]:                       # ties at the cut: same similarity values chosen
        a = np.sort(full[i][adj[i] > 0])
        b = np.sort(full[i][ref[i] > 0])
        assert np.array_equal(a, b), i
    return ora


def test_string_modality_oracle_matches_reference_fixture(golden_dir):
    for W, k, users, tags, au, at in _cases(golden_dir):
        assert np.array_equal(ops_oracle.username_adjacency(users), au)
        ora, _ = ops_oracle.tags_adjacency(tags, k)
        _check_tags(ora, at, tags, k)


@pytest.mark.gpu
def test_string_modalities_on_device(golden_dir):
    import torch
    from mused_b200 import matrix_operations as mo
    for W, k, users, tags, au, at in _cases(golden_dir):
        du = mo.create_adjacency_matrix(users.astype(object).reshape(-1, 1), "username", k_basis=k)
        assert np.array_equal(du, au)
        dt = mo.create_adjacency_matrix(tags.astype(object).reshape(-1, 1), "tags", k_basis=k)
        ora = _check_tags(dt, at, tags, k)
        assert np.array_equal(dt, ora)                # same tie rule as the oracle: identical everywhere
        # fused with a dense modality in one uint8 accumulator (main.py:44-56)
        rng = np.random.default_rng(W)
        F = rng.standard_normal((W, 24))
        fused = mo.fused_adjacency([F, users.reshape(-1, 1), tags.reshape(-1, 1)], ["dense", "username", "tags"], k_basis=k)
        want = (mo.create_adjacency_matrix(F, "dense", k_basis=k) > 0) | (au > 0) | (ora > 0)
        assert np.array_equal(fused.cpu().numpy() > 0, want)
    # tag LISTS (items instead of characters) and empty rows
    col = np.empty((5, 1), dtype=object)
    for i, t in enumerate([["sun", "sea"], ["sea", "sand"], "", ["sun"], ["rock"]]):
        col[i, 0] = t
    got = mo.create_adjacency_matrix(col, "tags", k_basis=1)
    assert got[0].tolist() == [0, 0, 0, 1, 0] and got[1].tolist() == [1, 0, 0, 0, 0] and got[2].sum() == 0
    assert got[3].tolist() == [1, 0, 0, 0, 0] and got[4].sum() == 1

"""GPU: the competitor sketches LM-FD and DI-FD (swfd/lmfd.py:18-117, swfd/difd.py:41-142) on the device merge
operator (mused_fd_merge) against outputs of the UNMODIFIED reference (tests/golden/competitors.npz, made by
oracle/gen_golden_competitors.py): B^T B of every query to 1e-9 of its largest entry, identical get_size."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["lmfd", "difd"])
def test_competitor_sketch_matches_reference(golden_dir, name):
    from mused_b200.swfd.lmfd import LMFD
    from mused_b200.swfd.difd import DIFD
    g = np.load(os.path.join(golden_dir, "competitors.npz"))
    A, N, ell, qstep = g["A"], int(g["N"]), int(g["ell"]), int(g["qstep"])
    sk = LMFD(N, A.shape[1], ell) if name == "lmfd" else DIFD(N, A.shape[1], ell, float(g["R"]))
    qi = 0
    for t in range(A.shape[0]):
        sk.fit(A[t:t + 1])
        if (t + 1) % qstep == 0:
            B = sk.get()[0]
            ref = g[name + "_gram"][qi]
            assert np.abs(B.T @ B - ref).max() <= 1e-9 * np.abs(ref).max()
            assert sk.get_size() == int(g[name + "_size"][qi])
            qi += 1
    assert qi == len(g[name + "_size"])


def test_fd_merge_against_lapack():
    import torch
    from mused_b200.swfd.lmfd import fd_merge
    rng = np.random.default_rng(0)
    for n, d, ell in ((40, 96, 12), (9, 64, 12), (200, 64, 16), (70, 64, 64)):
        M = rng.standard_normal((n, d)) * np.exp(rng.standard_normal((n, 1)))
        out = fd_merge(torch.as_tensor(M).cuda(), ell).cpu().numpy()
        s = np.linalg.svd(M, compute_uv=False) ** 2
        keep = s[:ell] - (s[ell] if len(s) > ell else 0.0)
        u, sv, vt = np.linalg.svd(M, full_matrices=False)
        ref = (vt[:len(keep)] * np.sqrt(np.maximum(keep, 0))[:, None])
        assert np.abs(out.T @ out - ref.T @ ref).max() <= 1e-9 * s[0]

"""Whole-GPU block Jacobi (csrc/eig_big.cuh, C ABI mused_sym_eig) against LAPACK."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _gram(n, d, rank, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((n, rank)) @ rng.standard_normal((rank, d)) * scale
    if rank >= min(n, d):
        A = rng.standard_normal((n, d)) * scale
    A *= np.exp(rng.standard_normal(n))[:, None]           # graded rows
    return A @ A.T


@pytest.mark.parametrize("n,d,rank", [(1, 4, 4), (2, 8, 8), (7, 16, 16), (33, 64, 64), (130, 96, 40),
                                      (257, 512, 512), (500, 300, 300), (1024, 1024, 1024),
                                      (1408, 1024, 1024)])
def test_sym_eig_matches_lapack(n, d, rank):
    from mused_b200.linalg import sym_eig
    G = _gram(n, d, rank, seed=n)
    ev, V, sweeps = sym_eig(torch.as_tensor(G).cuda(), return_sweeps=True)
    ev, V = ev.cpu().numpy(), V.cpu().numpy()
    ref = np.linalg.eigvalsh(G)[::-1]
    lmax = ref[0]
    assert 1 <= sweeps < 40
    assert np.all(np.diff(ev) <= 0)
    # eigenvalues that matter to a sketch (sigma >= 1e-3 sigma_max) to full accuracy; the columns of
    # the null space stop at the convergence threshold
    big = ref >= 1e-6 * lmax
    assert np.max(np.abs(ev - ref)[big]) <= 1e-13 * lmax * max(1, n / 64)
    assert np.max(np.abs(ev - ref)) <= 1e-8 * lmax
    k = min(n, d, rank)
    Vb = V[big]
    # orthonormal rows and G v = lambda v on the numerically significant part of the spectrum
    assert np.max(np.abs(Vb @ Vb.T - np.eye(Vb.shape[0]))) < 1e-7
    assert np.max(np.abs(Vb @ G - ev[big][:, None] * Vb)) < 1e-7 * lmax
    assert big.sum() <= k


def test_sym_eig_zero_and_diagonal():
    from mused_b200.linalg import sym_eig
    ev, V = sym_eig(torch.zeros((10, 10), dtype=torch.float64, device="cuda"))
    assert torch.all(ev == 0)
    D = torch.diag(torch.arange(1, 41, dtype=torch.float64, device="cuda"))
    ev, V, sweeps = sym_eig(D, return_sweeps=True)
    assert sweeps == 1
    assert torch.equal(ev, torch.arange(40, 0, -1, dtype=torch.float64, device="cuda"))
    assert torch.allclose(V.abs(), torch.flip(torch.eye(40, dtype=torch.float64, device="cuda"), [0]))

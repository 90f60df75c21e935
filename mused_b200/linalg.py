"""Symmetric eigensolve of a Gram matrix on the B200 (csrc/eig_big.cuh).

The reference takes the thin SVD of the stacked rows with LAPACK gesvd
(swfd/sliding_windows_frequent_directions.py:160); the device path forms the
float64 Gram of the stack and calls this solver: singular values =
sqrt(eigenvalues), left singular vectors = eigenvectors.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def sym_eig(G, *, return_sweeps=False):
    """Eigen-decomposition of the symmetric PSD matrix ``G`` (n x n, float64).

    Returns ``(evals, vecs)``: eigenvalues in DESCENDING order and the matching
    eigenvectors as the ROWS of ``vecs`` (torch tensors on G's device).
    """
    lib = _lib.load()
    if not isinstance(G, torch.Tensor):
        G = torch.as_tensor(G, dtype=torch.float64).to("cuda:0")
    if G.dim() != 2 or G.shape[0] != G.shape[1]:
        raise ValueError("G must be square")
    G = G.to(torch.float64).contiguous()
    n = G.shape[0]
    evals = torch.empty(n, dtype=torch.float64, device=G.device)
    vecs = torch.empty((n, n), dtype=torch.float64, device=G.device)
    sweeps = C.c_int32(0)
    with torch.cuda.device(G.device):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.mused_sym_eig(G.data_ptr(), n, n, evals.data_ptr(), vecs.data_ptr(), n,
                                     C.byref(sweeps), st))
    if return_sweeps:
        return evals, vecs, int(sweeps.value)
    return evals, vecs

"""LM-FD (logarithmic-method sliding-window Frequent Directions) on the device operators -- the competitor
sketch of /root/reference/swfd/lmfd.py:18-117 with the same ``fit / get / get_size`` protocol
(swfd/test_synthetic.py:53-54).

The block ladder (which blocks exist, when two of them merge, when a block expires) is host bookkeeping on
scalars exactly as in the reference; every block's rows live on the GPU as float64 tensors and the only
arithmetic -- the FD merge of two stacked blocks (lmfd.py:27-47: thin SVD, top-l cut, shrink by the (l+1)-th
squared singular value) -- is the C-ABI operator ``mused_fd_merge``.  A merged block is kept as l rows with
zero rows beyond its rank (the reference keeps only the non-padded rows; zero rows change no later singular
value, and ``get_size`` counts the reference's row numbers, which are tracked next to the tensors).
"""
from __future__ import annotations

from collections import deque

import numpy as np
import torch

from .. import _lib


def fd_merge(rows, sketch_dim):
    """Top-l FD shrink of a stack of rows (torch float64 CUDA, n x d) -> l x d tensor (mused_fd_merge)."""
    lib = _lib.load()
    rows = rows.contiguous()
    n, d = rows.shape
    out = torch.empty((sketch_dim, d), dtype=torch.float64, device=rows.device)
    s2 = torch.empty(sketch_dim, dtype=torch.float64, device=rows.device)
    with torch.cuda.device(rows.device):
        _lib.check(lib.mused_fd_merge(rows.data_ptr(), n, d, d, int(sketch_dim), out.data_ptr(), s2.data_ptr(), None,
                                      torch.cuda.current_stream().cuda_stream))
    return out


class _Block:
    __slots__ = ("rows", "n_rows", "start", "end", "size")

    def __init__(self, start, end):
        self.rows = None          # torch (n x d) float64 on the device, or None
        self.n_rows = 0           # rows the reference's array would have
        self.start, self.end, self.size = start, end, 0.0


class LMFD:
    def __init__(self, N, d, sketch_dim, *, device=0):
        self.N, self.d, self.sketch_dim = N, int(d), int(sketch_dim)
        self.b = self.sketch_dim                              # blocks per level before a merge (lmfd.py:25)
        self.L = 1
        self.levels = {1: deque()}
        self.time = np.uint64(0)
        self._dev = torch.device("cuda", int(device))
        self.B = _Block(self.time, self.time)

    # lmfd.py:27-47
    def _merge(self, a, a_rows, b, b_rows):
        if a is not None and b is not None:
            return fd_merge(torch.cat([a, b], dim=0), self.sketch_dim), min(a_rows + b_rows, self.sketch_dim, self.d)
        if a is not None:
            return a, a_rows
        if b is not None:
            return b, b_rows
        return torch.zeros((self.sketch_dim, self.d), dtype=torch.float64, device=self._dev), self.sketch_dim

    def _row(self, X):
        x = torch.as_tensor(np.asarray(X, dtype=np.float64) if not isinstance(X, torch.Tensor) else X)
        return x.to(device=self._dev, dtype=torch.float64).reshape(1, self.d)

    def fit(self, X, t=None):
        """One arriving row (1, d) (lmfd.py:50-94)."""
        self.time = np.uint64(t) if t is not None else self.time + np.uint64(1)
        # expiry, top level first; an emptied level above the first disappears (:56-66)
        for lv in sorted(self.levels, reverse=True):
            q = self.levels[lv]
            while q and q[0].end + self.N <= self.time:
                q.popleft()
            if not q and lv != 1:
                del self.levels[lv]
                self.L = lv - 1
        x = self._row(X)
        self.B.size += float((x * x).sum().item())
        self.B.end = self.time
        if self.B.rows is None:
            self.B.rows, self.B.start = x, self.time
        else:
            self.B.rows = torch.cat([self.B.rows, x], dim=0)
        self.B.n_rows += 1
        if self.B.size >= self.sketch_dim:                    # :77-79 the open block is full by ENERGY
            self.levels[1].append(self.B)
            self.B = _Block(self.time, self.time)
        for lv in range(1, self.L + 1):                       # :81-94
            if len(self.levels[lv]) >= self.b + 1:
                if lv + 1 not in self.levels:
                    self.levels[lv + 1] = deque()
                    self.L = lv + 1
                first = self.levels[lv].popleft()
                if first.size + self.levels[lv][0].size >= (2 ** (lv + 1)) * self.sketch_dim:
                    self.levels[lv + 1].append(first)
                else:
                    second = self.levels[lv].popleft()
                    blk = _Block(first.start, second.end)
                    blk.rows, blk.n_rows = self._merge(first.rows, first.n_rows, second.rows, second.n_rows)
                    blk.size = first.size + second.size
                    self.levels[lv + 1].append(blk)

    def get(self):
        """(sketch, None, None, None): the open block merged with every live block, top level first (:97-104)."""
        ret, n = (self.B.rows.clone(), self.B.n_rows) if self.B.rows is not None else (None, 0)
        for lv in range(self.L, 0, -1):
            for blk in self.levels[lv]:
                ret, n = self._merge(ret, n, blk.rows, blk.n_rows)
        return (ret[:n].cpu().numpy() if ret is not None else None), None, None, None

    def get_size(self):                                       # :106-115
        return self.B.n_rows + sum(blk.n_rows for q in self.levels.values() for blk in q)

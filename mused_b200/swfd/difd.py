"""DI-FD (dyadic-interval sliding-window Frequent Directions) on the device operators -- the competitor sketch
of /root/reference/swfd/difd.py:41-142 with the same ``fit / get / get_size`` protocol
(swfd/test_synthetic.py:55-56).

Level i keeps dyadic blocks; every block is a RobustFrequentDirections sketch of L * 2^i rows (capped at d) fed
in batches of that many rows (difd.py:20-39 -> frequent_directions.py:39-58, :86-93).  Block scheduling is host
bookkeeping on scalars as in the reference; the rows live on the GPU and every batch update and the final
merge of ``get`` are the C-ABI operator ``mused_fd_merge`` (thin SVD + top-l cut + shrink of the stacked rows).
"""
from __future__ import annotations

from collections import deque

import numpy as np
import torch

from .lmfd import fd_merge


class _DyadicBlock:
    """difd.py:11-38: an RFD sketch of `dim` rows plus the rows that arrived since its last batch update."""

    def __init__(self, d, dim, device, start=1, end=0):
        self.d, self.dim, self.device = d, dim, device
        self.sketch = torch.zeros((dim, d), dtype=torch.float64, device=device)
        self.pending = []
        self.count, self.start, self.end, self.size = 0, start, end, 0.0

    def append(self, x):
        self.count += 1
        self.pending.append(x)
        if len(self.pending) >= self.dim:
            self.flush()

    def flush(self):
        # RobustFrequentDirections.fit(buffer, batch_size=dim): one shrink of [sketch; batch] per batch of dim rows
        while self.pending:
            batch, self.pending = self.pending[:self.dim], self.pending[self.dim:]
            self.sketch = fd_merge(torch.cat([self.sketch] + batch, dim=0), self.dim)

    def get_sketch(self):
        self.flush()
        return self.sketch


class DIFD:
    def __init__(self, N, d, sketch_dim, R, *, device=0):
        self.N, self.d, self.sketch_dim, self.R = N, int(d), int(sketch_dim), R
        self.L = int(np.ceil(np.log2(R * sketch_dim)))                     # difd.py:48-49
        self._dev = torch.device("cuda", int(device))
        self.levels = {i: deque() for i in range(1, self.L + 1)}
        self.Bs = {i: self._fresh(i) for i in range(1, self.L + 1)}
        self.time = 0
        self.base_len = 0

    def _fresh(self, i, start=1, end=0):
        return _DyadicBlock(self.d, min(self.d, self.L * (2 ** i)), self._dev, start, end)   # :53-54, :87-88

    def fit(self, X):
        """One arriving row (1, d) (difd.py:58-88)."""
        self.time += 1
        x = torch.as_tensor(np.asarray(X, dtype=np.float64) if not isinstance(X, torch.Tensor) else X)
        x = x.to(device=self._dev, dtype=torch.float64).reshape(1, self.d)
        e = float((x * x).sum().item())
        for lv in range(1, self.L + 1):
            q = self.levels[lv]
            while q and q[0].start < self.time - self.N + 1:               # blocks that left the window
                q.popleft()
            blk = self.Bs[lv]
            blk.size += e
            blk.end = self.time
            blk.append(x)
        if self.Bs[1].size > self.N / self.sketch_dim:                      # :73-88 close dyadic blocks
            self.base_len += 1
            v = min((self.base_len & -self.base_len).bit_length(), self.L + 1)
            for i in range(1, v):
                if self.Bs[i].start >= self.time - self.N + 1:
                    self.Bs[i].flush()
                    self.levels[i].append(self.Bs[i])
                self.Bs[i] = self._fresh(i, start=self.time + 1, end=self.time)

    def get(self):
        """(sketch, None, None, None): the blocks that tile the window, merged (difd.py:90-129)."""
        a = self.time
        b = max(1, int(self.time - self.N + 1))
        parts = [self.Bs[1].get_sketch()]
        for i in range(self.L, 0, -1):
            q = self.levels[i]
            if not q:
                continue
            lb = q[0]
            if lb.start >= self.time - self.N + 1:
                if lb.end <= a:
                    parts.append(lb.get_sketch())
                    a = lb.start
                if len(q) != 1:
                    rb = q[-1]
                    if rb.start >= b:
                        parts.append(rb.get_sketch())
                        b = rb.end
                else:
                    b = lb.end
        stack = torch.cat(parts, dim=0)
        out = fd_merge(stack, self.sketch_dim)
        return out[:min(stack.shape[0], self.sketch_dim, self.d)].cpu().numpy(), None, None, None

    def get_size(self):                                                     # :131-139
        return sum(min(blk.count, self.sketch_dim) for q in self.levels.values() for blk in q)

"""NumPy shard engine with the phase protocol of mused_b200/sharded.py -- TEST
INFRASTRUCTURE ONLY (see oracle/__init__.py).

Lets the CPU suite run the product's column-sharded DRIVER
(mused_b200.sharded.ColumnShardedSWFD: chunking, exchange order, column slicing,
gather) over the gloo backend with world_size 2, with this float64 engine standing
in for the CUDA one.  The arithmetic is the batched schedule of
oracle/swfd_batched_oracle.py taken through the Gram route (the only route that
shards: G = sum_r R_r R_r^T, eigensolve replicated, rotate local), i.e. exactly what
csrc/bswfd.cu does per rank; reference lines as in swfd_batched_oracle.py.
"""
from __future__ import annotations

from collections import deque

import numpy as np
import torch


class _Side:
    def __init__(self, d, ell, last_dump):
        self.sketch = np.zeros((ell, d))
        self.s2 = np.zeros(ell)
        self.delta = 0.0
        self.last_dump = int(last_dump)
        self.queue = deque()
        self.n_flush = self.n_dump = self.n_raw = 0


class ShardEngineOracle:
    def __init__(self, N, R, d_local, sketch_dim, beta=1.0, batch=None, **kw):
        self.N, self.d, self.ell = int(N), int(d_local), int(sketch_dim)
        self.batch = int(batch) if batch else self.ell
        self.logR = int(np.ceil(np.log2(R)))
        self.thr = [float((2 ** j) * self.N / self.ell) for j in range(self.logR)]
        self.fd = [_Side(self.d, self.ell, 1) for _ in range(self.logR)]
        self.aux = [_Side(self.d, self.ell, 1) for _ in range(self.logR)]
        self.period = 0
        self.time = 0
        self.done = 0
        self.X = np.zeros((self.ell, self.d))
        self.pend = 0
        self.cap = (2 + 8 / beta) * self.ell
        S, n2 = 2 * self.logR, 2 * self.ell
        self.norm2 = torch.zeros(self.ell, dtype=torch.float64)
        self.gram = torch.zeros(max(S, 1) * n2 * n2, dtype=torch.float64)
        self.qmax = self.ell + int(self.cap) + self.batch + self.ell + 4
        self.query_gram = torch.zeros(self.qmax * self.qmax, dtype=torch.float64)
        self._stacks = None

    # ---- phase protocol -----------------------------------------------------------
    def room(self):
        t_a = self.done + 1
        to_b = self.N - (t_a % self.N) if t_a % self.N else self.N
        return min(self.batch, to_b) - self.pend

    def pending(self):
        return self.pend

    def stage(self, X):
        X = np.asarray(X.cpu() if hasattr(X, "cpu") else X, dtype=np.float64)
        m = X.shape[0]
        assert m <= self.room()
        self.X[self.pend:self.pend + m] = X
        self.norm2[self.pend:self.pend + m] = torch.from_numpy(np.einsum("ij,ij->i", X, X))
        self.pend += m
        self.time += m

    def _sides(self):
        for j in range(self.logR):
            yield j, self.fd[j]
            yield j, self.aux[j]

    def chunk_gram(self):
        n, t_a = self.pend, self.done + 1
        t_b = t_a + n - 1
        while self.period < t_a // self.N:
            for j in range(self.logR):
                self.fd[j] = self.aux[j]
                self.aux[j] = _Side(self.d, self.ell, t_a)
            self.period += 1
        norm2 = self.norm2.numpy()[:n]                 # already summed over the shards
        X = self.X[:n]
        n2 = 2 * self.ell
        G = self.gram.numpy().reshape(-1, n2, n2)
        G[:] = 0.0
        self._stacks = []
        for i, (j, side) in enumerate(self._sides()):
            q = side.queue
            while len(q) != 0 and (len(q) > self.cap or q[0][2] + self.N <= t_b):
                q.popleft()
            raw = norm2 >= self.thr[j]
            for r in np.where(raw)[0]:
                q.append((X[r:r + 1].copy(), side.last_dump, t_a + int(r)))
                side.last_dump = t_a + int(r) + 1
                side.n_raw += 1
            Xp = X[~raw]
            stack = np.vstack([side.sketch, Xp])
            self._stacks.append(stack)
            m = stack.shape[0]
            if len(Xp):
                G[i, :m, :m] = stack @ stack.T

    def chunk_finish(self):
        n, t_b = self.pend, self.done + self.pend
        n2, ell = 2 * self.ell, self.ell
        G = self.gram.numpy().reshape(-1, n2, n2)
        for i, (j, side) in enumerate(self._sides()):
            stack = self._stacks[i]
            m = stack.shape[0]
            if m == ell:                                # no pending rows for this instance
                continue
            lam, U = np.linalg.eigh(G[i, :m, :m])
            order = np.argsort(-lam, kind="stable")
            lam, U = np.maximum(lam[order], 0.0), U[:, order]
            cut = lam[ell]
            side.delta += cut / 2.0
            s2 = np.maximum(lam[:ell] - cut, 0.0)
            coef = np.where(lam[:ell] > 0, np.sqrt(s2 / np.where(lam[:ell] > 0, lam[:ell], 1.0)), 0.0)
            side.sketch = (U[:, :ell] * coef).T @ stack
            side.s2 = s2
            side.n_flush += 1
            idx = np.where(s2 >= self.thr[j])[0]
            if len(idx):
                s0 = min(side.last_dump, t_b)
                for k in idx:
                    side.queue.append((side.sketch[k:k + 1].copy(), s0, t_b))
                    side.sketch[k, :] = 0.0
                    side.s2[k] = 0.0
                    side.n_dump += 1
                side.last_dump = t_b + 1
        self.done += n
        self.pend = 0

    def get_begin(self):
        assert self.pend == 0
        rj = 0
        for j in range(self.logR):
            q = self.fd[j].queue
            if len(q) == 0:
                break
            if self.time - q[0][1] >= min(self.N - 1, self.time - 1):
                rj = j
                break
        self._q = self.fd[rj]
        side = self._q
        if len(side.queue) == 0:
            self._qstack = None
            return rj, 0
        self._qstack = np.vstack([side.sketch, *(e[0] for e in side.queue)])
        ng = self._qstack.shape[0]
        QG = self.query_gram.numpy()
        QG[:] = 0.0
        QG[:ng * ng] = (self._qstack @ self._qstack.T).reshape(-1)
        return rj, ng

    def get_finish(self):
        side, ell = self._q, self.ell
        if self._qstack is None:
            s2 = side.s2
            vt = np.where(s2[:, None] > 0, side.sketch / np.sqrt(np.where(s2 > 0, s2, 1.0))[:, None], 0.0)
            return side.sketch.copy(), s2.copy(), vt, side.delta
        M = self._qstack
        ng = M.shape[0]
        lam, U = np.linalg.eigh(self.query_gram.numpy()[:ng * ng].reshape(ng, ng))
        order = np.argsort(-lam, kind="stable")
        lam, U = np.maximum(lam[order], 0.0), U[:, order]
        cut = lam[ell] if ng > ell else 0.0
        top = lam[:ell]
        vt = (U[:, :ell] / np.sqrt(np.where(top > 0, top, 1.0))).T @ M
        vt[top <= 0] = 0.0
        s = np.maximum(top - cut, 0.0)
        return vt * np.sqrt(s)[:, None], s, vt, 0.0

    def get_size(self):
        return sum(len(self.fd[j].queue) + len(self.aux[j].queue) + 2 * self.ell for j in range(self.logR))

    def debug_state(self):
        return [dict(level=j, delta=float(s.delta), s2=np.array(s.s2), last_dump_time=int(s.last_dump),
                     q_s=[int(e[1]) for e in s.queue], q_t=[int(e[2]) for e in s.queue],
                     n_flush=s.n_flush, n_dump=s.n_dump, n_raw=s.n_raw) for j, s in self._sides()]

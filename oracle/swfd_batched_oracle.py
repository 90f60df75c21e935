"""Float64 NumPy statement of the BATCHED flush schedule of the SWFD path.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference flushes its row buffer when an energy bound crosses the level
threshold (swfd/frequent_directions.py:122-125) -- a per-row dependency chain.
BASELINE.json's north_star names the batched form instead ("row-batch append
into the 2l x d sketch buffer", one Gram/eig/rotate per batch); this file states
that schedule exactly, reusing the reference's own arithmetic for every step:

  * shrink of [sketch; batch]      swfd/frequent_directions.py:39-58 (fit with a
                                   batch) + :86-93 (delta += s2[l]/2, s2 -= s2[l])
  * dump of EVERY direction with   swfd/fastfdwithdump.py:115-131 (the reference's
    sigma^2 >= threshold           own multi-direction dump) instead of :127-142
  * raw-row snapshots, periods,    swfd/sliding_windows_frequent_directions.py:56-138
    expiry, (s, t) stamps
  * level ladder, queue cap,       swfd/seq_based_swfd.py:11-103
    routing, level choice, get()

Schedule (the only departure from the reference): the stream is cut into CHUNKS;
a chunk ends after `batch` rows (default l), before a row whose time is a
multiple of N (period boundary), or at a get().  Per chunk and per FD instance:
period swap (chunk start), queue trim against the chunk's last time, raw rows
(||x||^2 >= threshold) become snapshots with their own times, the other rows are
shrunk in ONE batch, then every direction with sigma^2 >= threshold is dumped
with stamp (s = min(last_dump, t_end), t = t_end).  The result does not depend
on how the caller slices the rows over fit() calls.

This schedule is not pinned by the reference's outputs (its decisions differ by
construction); what is pinned is (a) its covariance error against the
reference's on the same stream (tests/test_batched_oracle.py) and (b) the CUDA
engine against THIS file.  The reference-schedule engine stays the strict
parity path.
"""
from __future__ import annotations

from collections import deque

import numpy as np
from scipy import linalg


def _thin_svd(M):
    _, s, vt = linalg.svd(M, full_matrices=False, lapack_driver="gesvd")
    return s, vt


class _Side:
    def __init__(self, d, ell, last_dump):
        self.sketch = np.zeros((ell, d))
        self.s2 = np.zeros(ell)
        self.delta = 0.0
        self.last_dump = int(last_dump)
        self.queue = deque()          # entries (v 1 x d, s, t), one row each
        self.n_flush = 0
        self.n_dump = 0
        self.n_raw = 0


class BatchedSeqSWFD:
    def __init__(self, N, R, d, sketch_dim, beta=1.0, batch=None, **kwargs):
        self.N, self.d, self.ell, self.beta = int(N), int(d), int(sketch_dim), beta
        self.batch = int(batch) if batch else self.ell
        assert 1 <= self.batch <= self.ell
        self.logR = int(np.ceil(np.log2(R)))                       # seq_based_swfd.py:25
        if "upper_F_norm" in kwargs:                               # :31-37
            self.logR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / self.N)))
        self.thresholds = [float((2 ** j) * self.N / self.ell) for j in range(self.logR)]
        self.fd = [_Side(d, self.ell, 1) for _ in range(self.logR)]
        self.aux = [_Side(d, self.ell, 1) for _ in range(self.logR)]
        self.period = 0
        self.time = 0                  # rows handed to fit()
        self.done = 0                  # rows already folded into the state
        self.pend = []                 # rows [done, time)
        self.cap = (2 + 8 / beta) * self.ell                       # :61

    # ---- one FD instance ----------------------------------------------------
    def _shrink(self, side, Xp):
        ell = self.ell
        stack = np.vstack([side.sketch, Xp])
        s, vt = _thin_svd(stack)
        s2 = s ** 2
        if len(s2) > ell:                                          # frequent_directions.py:87-92
            side.delta += s2[ell] / 2.0
            s2 = s2[:ell] - s2[ell]
            vt = vt[:ell]
        side.s2 = s2
        side.sketch = vt * np.sqrt(s2).reshape(-1, 1)
        side.n_flush += 1

    def _dump(self, side, thr, t_end):
        idx = np.where(side.s2 >= thr)[0]                          # fastfdwithdump.py:116
        if len(idx) == 0:
            return
        s0 = min(side.last_dump, t_end)
        for i in idx:
            side.queue.append((side.sketch[i:i + 1].copy(), s0, t_end))
            side.sketch[i, :] = 0.0
            side.s2[i] = 0.0
            side.n_dump += 1
        side.last_dump = t_end + 1

    # ---- one chunk ------------------------------------------------------------
    def _chunk(self, X, t_a):
        n = X.shape[0]
        t_b = t_a + n - 1
        N = self.N
        while self.period < t_a // N:                              # sliding_windows...py:99-107
            for j in range(self.logR):
                self.fd[j] = self.aux[j]
                self.aux[j] = _Side(self.d, self.ell, t_a)
            self.period += 1
        norm2 = np.einsum("ij,ij->i", X, X)
        for j in range(self.logR):
            thr = self.thresholds[j]
            raw = norm2 >= thr                                     # seq_based_swfd.py:76-77
            Xp = X[~raw]
            for side in (self.fd[j], self.aux[j]):
                q = side.queue
                while len(q) != 0 and (len(q) > self.cap or q[0][2] + N <= t_b):   # :59-74
                    q.popleft()
                for r in np.where(raw)[0]:                         # sliding_windows...py:56-80
                    t_r = t_a + int(r)
                    q.append((X[r:r + 1].copy(), side.last_dump, t_r))
                    side.last_dump = t_r + 1
                    side.n_raw += 1
                if len(Xp):
                    self._shrink(side, Xp)
                    self._dump(side, thr, t_b)

    def _drain(self, force):
        while self.pend:
            t_a = self.done + 1
            room = self.N - (t_a % self.N) if t_a % self.N else self.N   # rows until the next multiple of N
            n = min(self.batch, room, len(self.pend))
            full = (n == self.batch) or (n == room)
            if not full and not force:
                return
            X = np.vstack(self.pend[:n])
            del self.pend[:n]
            self._chunk(X, t_a)
            self.done += n

    # ---- public surface (seq_based_swfd.py:47-103) ------------------------------
    def fit(self, X):
        X = np.asarray(X, dtype=np.float64)
        if X.ndim == 1:
            X = X.reshape(1, -1)
        for r in range(X.shape[0]):
            self.pend.append(X[r:r + 1])
        self.time += X.shape[0]
        self._drain(False)

    def pick_level(self):                                          # :83-98
        rj = 0
        for j in range(self.logR):
            q = self.fd[j].queue
            if len(q) == 0:
                break
            if self.time - q[0][1] >= min(self.N - 1, self.time - 1):
                rj = j
                break
        return rj

    def get(self):                                                 # :100 + sliding_windows...py:154-173
        if self.logR <= 0:
            raise KeyError(0)
        self._drain(True)
        side = self.fd[self.pick_level()]
        ell = self.ell
        if len(side.queue) != 0:
            stack = np.vstack([side.sketch, *(e[0] for e in side.queue)])
            s, vt = _thin_svd(stack)
            s = s ** 2
            if s.shape[0] > ell:
                s = s[:ell] - s[ell]
            vt = vt[:ell]
            return vt * np.sqrt(s).reshape(-1, 1), s, vt, 0.0
        s2 = side.s2
        vt = np.where(s2[:, None] > 0, side.sketch / np.sqrt(np.where(s2 > 0, s2, 1.0))[:, None], 0.0)
        return side.sketch, s2, vt, side.delta

    def get_size(self):                                            # :102-103
        return sum(len(self.fd[j].queue) + len(self.aux[j].queue) + 2 * self.ell for j in range(self.logR))

    def debug_state(self):
        out = []
        for j in range(self.logR):
            for side in (self.fd[j], self.aux[j]):
                out.append(dict(level=j, delta=float(side.delta), s2=np.array(side.s2),
                                last_dump_time=int(side.last_dump),
                                q_s=[int(e[1]) for e in side.queue], q_t=[int(e[2]) for e in side.queue],
                                n_flush=side.n_flush, n_dump=side.n_dump, n_raw=side.n_raw))
        return out

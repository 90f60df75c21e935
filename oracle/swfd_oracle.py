"""NumPy/SciPy restatement of the reference's sliding-window Frequent
Directions path.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows, with quirks included:
  * swfd/frequent_directions.py:39-58   FD shrink through a thin LAPACK gesvd
  * swfd/frequent_directions.py:86-93   RFD delta += s2[l]/2, s2[:l] -= s2[l]
  * swfd/frequent_directions.py:96-153  buffer + energy-triggered flush, dump
  * swfd/sliding_windows_frequent_directions.py:27-197  one DS-FD level
  * swfd/seq_based_swfd.py:11-103       multi-level sequence-based SWFD

The state is laid out as plain records + lists (not the reference's class
tree) but every arithmetic statement is evaluated in the same order on the
same float64 values, so outputs are bit-identical to the reference on the
same LAPACK (asserted by oracle/gen_golden.py).
"""
from __future__ import annotations

import math
from collections import deque

import numpy as np
from scipy import linalg


def _thin_svd(M):
    # frequent_directions.py:55 -- driver is pinned to gesvd upstream
    _, s, vt = linalg.svd(M, full_matrices=False, lapack_driver="gesvd")
    return s, vt


class FDDump:
    """frequent_directions.py:96-153 (FrequentDirectionsWithDump) on top of
    :21-64 (FrequentDirections) and :79-93 (RobustFrequentDirections)."""

    def __init__(self, d, sketch_dim, error):
        self.d = d
        self.ell = min(sketch_dim, d)                     # :98
        self.error = error
        self.sketch = np.zeros((self.ell, d))
        self.vt = np.zeros((self.ell, d))
        self.s2 = np.zeros(self.ell)
        self.delta = 0.0
        self.energy = 0.0
        self.pending = None                               # row buffer (:100)

    # -- :39-58 with batch_size = rows(buffer) (<= ell), one batch only
    def _shrink(self, X, batch):
        n = X.shape[0]
        for i in range(0, n, batch):
            stack = np.zeros((self.ell + batch, self.d))
            blk = X[i:i + batch, :]
            stack[:self.ell, :] = self.sketch
            stack[self.ell:self.ell + blk.shape[0], :] = blk
            s, self.vt = _thin_svd(stack)
            self.s2 = s ** 2
            if len(self.s2) > self.ell:                   # :87
                self.delta += self.s2[self.ell] / 2.0
                self.s2 = self.s2[:self.ell] - self.s2[self.ell]
                self.vt = self.vt[:self.ell]
            self.sketch = self.vt * np.sqrt(self.s2).reshape(-1, 1)

    def flush(self):                                      # :103-109
        if self.pending is not None:
            self._shrink(self.pending, min(self.pending.shape[0], self.ell))
            self.energy = self.s2[0]
            self.pending = None

    def fit(self, X):                                     # :115-125
        self.energy = self.energy + X @ X.T
        self.pending = X if self.pending is None else np.concatenate([self.pending, X])
        if len(self.pending) >= self.ell:
            self.flush()
        elif self.energy >= self.error:
            self.flush()

    def dump(self):                                       # :127-142
        if self.s2[0] >= self.error:
            v = np.sqrt(self.s2[0]) * self.vt[0:1]
            self.sketch[0, :] = 0
            self.s2[0] = 0
            self.vt[0, :] = 0
            # the three np.roll results are discarded upstream: rows stay put
            self.energy = self.s2[0]
            return v
        return None

    def get(self):                                        # :144-146, :60-61
        self.flush()
        return self.sketch, self.s2, self.vt, self.delta

    def get_sketch(self):                                 # :148-150
        self.flush()
        return self.sketch


class RFD:
    """frequent_directions.py:79-93 RobustFrequentDirections(d, sketch_dim)
    with the base-class fit(X, batch_size) of :39-58."""

    def __init__(self, d, sketch_dim=8):
        self.core = FDDump(d, sketch_dim, math.inf)
        self.core.ell = sketch_dim                        # base class does not clip to d (:32-36)
        self.core.sketch = np.zeros((sketch_dim, d))
        self.core.vt = np.zeros((sketch_dim, d))
        self.core.s2 = np.zeros(sketch_dim)

    def fit(self, X, batch_size=1):
        self.core._shrink(X, batch_size)

    def get(self):
        c = self.core
        return c.sketch, c.s2, c.vt, c.delta

    def get_sketch(self):
        return self.core.sketch


class _Side:
    """sliding_windows_frequent_directions.py:19-24 DumpFDwithSnapshotQueue;
    queue entries are (v, s, t) = FDSnapshot (:12-16)."""

    def __init__(self, d, ell, error, last_dump_time=np.uint64(1)):
        self.C = FDDump(d, ell, error)
        self.queue = deque()
        self.last_dump_time = last_dump_time


class Level:
    """sliding_windows_frequent_directions.py:27-197 SlidingWindowFD (C=1)."""

    def __init__(self, N, d, ell, error=None):
        self.error = N * 1.0 / ell if error is None else error     # :37-40
        self.fd = _Side(d, ell, self.error)
        self.aux = _Side(d, ell, self.error)
        self.period = 0
        self.N = np.uint64(N)
        self.d, self.ell = d, ell
        self.time = np.uint64(0)

    def _tick(self, t=None):                              # :60-75 / :90-107
        if t is not None:                                 # explicit arrival time (TimeBasedSWFD)
            self.time = np.uint64(t)
        else:
            self.time += np.uint64(1)
        while self.period < self.time // self.N:
            self.fd = self.aux
            self.aux = _Side(self.d, self.ell, self.error, last_dump_time=self.time)
            self.period += 1

    def _push(self, side, v):                             # :73-80, :127-138
        s = side.last_dump_time
        side.last_dump_time = self.time + 1
        side.queue.append((v, s, self.time))

    def append(self, X, t=None):                          # :56-80 raw-row snapshot
        self._tick(t)
        self._push(self.fd, X)
        self._push(self.aux, X)

    def fit(self, X, t=None):                             # :83-138
        self._tick(t)
        for side in (self.fd, self.aux):                  # :109-121
            q = side.queue
            while len(q) != 0 and q[0][2] + self.N <= self.time:
                q.popleft()
        for side in (self.fd, self.aux):                  # :124-138
            side.C.fit(X)
            v = side.C.dump()
            if v is not None:
                self._push(side, v)

    def get(self):                                        # :154-173
        q = self.fd.queue
        C = self.fd.C.get_sketch()
        if len(q) != 0:
            stack = np.vstack([C, *(e[0] for e in q)])
            s, vt = _thin_svd(stack)
            s = s ** 2
            if s.shape[0] > self.ell:
                s = s[:self.ell] - s[self.ell]
            vt = vt[:self.ell]
            return vt * np.sqrt(s).reshape(-1, 1), s, vt, 0.0
        return self.fd.C.get()

    def get_size(self):                                   # :196-197
        return len(self.fd.queue) + len(self.aux.queue) + 2 * self.ell


class SeqSWFD:
    """seq_based_swfd.py:11-103 SeqBasedSWFD(N, R, d, sketch_dim, beta, **kw)."""

    def __init__(self, N, R, d, sketch_dim, beta=1.0, **kwargs):
        self.N = np.uint64(N)
        self.d, self.R, self.beta, self.ell = d, R, beta, sketch_dim
        self.logR = int(np.ceil(np.log2(R)))              # :25
        self.levels = {}
        if "upper_F_norm" in kwargs:                      # :31-37
            self.logR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / self.N)))
            base = self.N / self.ell
            for j in range(self.logR):
                self.levels[j] = Level(self.N, d, self.ell, error=base * (2 ** j))
        else:                                             # :39-42
            for j in range(self.logR):
                self.levels[j] = Level(self.N, d, self.ell, error=(2 ** j) * self.N / self.ell)
        self.time = np.uint64(0)

    def fit(self, X):                                     # :47-80
        self.time += np.uint64(1)
        cap = (2 + 8 / self.beta) * self.ell              # :61
        for j in range(self.logR):
            lv = self.levels[j]
            for side in (lv.fd, lv.aux):
                q = side.queue
                while len(q) != 0 and (len(q) > cap or q[0][2] + self.N <= self.time):
                    q.popleft()
            if X @ X.T >= lv.aux.C.error:                 # :76-77
                lv.append(X)
            else:
                lv.fit(X)

    def pick_level(self):                                 # :83-98
        j = rj = 0
        while j < self.logR:
            q = self.levels[j].fd.queue
            if len(q) == 0:
                break
            if self.time - q[0][1] >= min(self.N - 1, self.time - 1):
                rj = j
                break
            j += 1
        return rj

    def get(self):                                        # :100
        return self.levels[self.pick_level()].get()

    def get_size(self):                                   # :102-103
        return sum(self.levels[j].get_size() for j in self.levels)

    # ---- introspection used only by the differential tests -----------------
    def debug_state(self):
        out = []
        for j in range(self.logR):
            lv = self.levels[j]
            for side in (lv.fd, lv.aux):
                out.append(dict(
                    level=j, energy=float(np.asarray(side.C.energy).reshape(-1)[0]) if np.ndim(side.C.energy) else float(side.C.energy),
                    delta=float(side.C.delta), s2=np.array(side.C.s2, dtype=float),
                    pending=0 if side.C.pending is None else len(side.C.pending),
                    last_dump_time=int(side.last_dump_time),
                    q_s=[int(e[1]) for e in side.queue], q_t=[int(e[2]) for e in side.queue]))
        return out


class TimeSWFD(SeqSWFD):
    """seq_based_swfd.py:106-207 TimeBasedSWFD(N, R, d, sketch_dim, beta=4, **kw): rows carry an
    explicit arrival time, the ladder has ceil(log2(N R / l)) levels with thresholds 2^j."""

    def __init__(self, N, R, d, sketch_dim, beta=4.0, **kwargs):
        self.N = np.uint64(N)
        self.d, self.R, self.beta, self.ell = d, R, beta, sketch_dim
        self.logR = int(np.ceil(np.log2(N * R / sketch_dim)))                      # :120
        self.levels = {}
        if "upper_F_norm" in kwargs:                                               # :131-141
            self.logR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / self.ell)))
            base = 1.0
            if "lower_F_norm" in kwargs:
                self.logR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / kwargs["lower_F_norm"])))
                base = kwargs["lower_F_norm"] / self.ell
            for j in range(self.logR):
                self.levels[j] = Level(self.N, d, self.ell, error=base * (2 ** j))
        else:                                                                      # :143-146
            for j in range(self.logR):
                self.levels[j] = Level(self.N, d, self.ell, error=2 ** j)
        self.time = np.uint64(0)

    def fit(self, X, t):                                                           # :150-189
        self.time = t
        cap = (2 + 8 / self.beta) * self.ell
        for j in range(self.logR):
            lv = self.levels[j]
            for side in (lv.fd, lv.aux):
                q = side.queue
                while len(q) != 0 and (len(q) > cap or q[0][2] + self.N <= self.time):
                    q.popleft()
            if X @ X.T >= lv.aux.C.error:                                          # :185-186
                lv.append(X, t)
            else:
                lv.fit(X, t)


def cov_error(A_w, B):
    """swfd/test_synthetic.py:83-90 -- ||A_w^T A_w - B^T B||_2 / ||A_w||_F^2."""
    A_f = linalg.norm(A_w) ** 2
    return linalg.norm(A_w.T @ A_w - B.T @ B, 2) / A_f

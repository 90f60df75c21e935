"""The reference's "Fast" sketch family on the device operators:

  * ``FastFrequentDirectionsWithDump``  swfd/fastfdwithdump.py:7-138 -- rows are stacked under the sketch, the Gram
    ``cov = sketch sketch^T`` is bordered row by row, its eigen-decomposition is taken only while the energy bound
    is above the dump threshold (every direction with sigma^2 >= threshold is deflated and returned at once), and
    a full FD shrink runs when the stack reaches 2 l rows;
  * ``FastSlidingWindowFD``             swfd/sliding_windows_frequent_directions.py:200-352 -- one DS-FD level on it
    (snapshots may hold several rows);
  * ``FastTimeBasedSWFD``               swfd/seq_based_swfd.py:209-311 -- the time-based level ladder.

Control flow and scalars (energy bound, periods, queues, stamps) are host Python exactly as in the reference,
which is itself a per-row Python loop; the state (stack, Gram, snapshots) stays on the GPU as float64 tensors and
the two decompositions are C-ABI operators: ``mused_sym_eig`` for the Gram (the reference's SVD of the symmetric
``cov``, :39) and ``mused_fd_merge`` for the full shrink (:98-108) and the query (:320-327).  Row-sized products
(bordering the Gram, deflation) are torch calls on the device.  Singular vectors are sign-ambiguous: results are
compared through B^T B (tests/test_fast_swfd_gpu.py).
"""
from __future__ import annotations

import ctypes as C
from collections import deque

import numpy as np
import torch

from .. import _lib
from .lmfd import fd_merge


def _sym_eig(G):
    """(evals descending, vecs with ROW i = eigenvector i) of a symmetric PSD device matrix (mused_sym_eig)."""
    lib = _lib.load()
    n = G.shape[0]
    G = G.contiguous()
    ev = torch.empty(n, dtype=torch.float64, device=G.device)
    vec = torch.empty((n, n), dtype=torch.float64, device=G.device)
    with torch.cuda.device(G.device):
        _lib.check(lib.mused_sym_eig(G.data_ptr(), n, n, ev.data_ptr(), vec.data_ptr(), n, None,
                                     torch.cuda.current_stream().cuda_stream))
    return ev, vec


class FastFrequentDirectionsWithDump:
    def __init__(self, d, sketch_dim, error, approx=False, *, device=0):
        self.d, self.sketch_dim = int(d), int(sketch_dim)
        self._dev = torch.device("cuda", int(device))
        self.sketch = torch.zeros((self.sketch_dim, self.d), dtype=torch.float64, device=self._dev)
        self.sigma_squared = torch.zeros(self.sketch_dim, dtype=torch.float64, device=self._dev)
        self.max_energy = 0.0
        self.error = float(error)
        self.cov = torch.zeros((self.sketch_dim, self.sketch_dim), dtype=torch.float64, device=self._dev)
        self.approx = approx

    def get_error(self):
        return self.error

    def _row(self, X):
        x = torch.as_tensor(np.asarray(X, dtype=np.float64)) if not isinstance(X, torch.Tensor) else X
        return x.to(device=self._dev, dtype=torch.float64).reshape(1, self.d)

    # fastfdwithdump.py:21-83: border the Gram with the newest row, deflate what crossed the threshold
    def _border_and_deflate(self):
        top, last = self.sketch[:-1], self.sketch[-1:]
        col = top @ last.T
        self.cov = torch.cat([torch.cat([self.cov, col], dim=1), torch.cat([col.T, last @ last.T], dim=1)], dim=0)
        if self.max_energy < self.error:
            return None
        lam, vec = _sym_eig(self.cov)                      # cov is symmetric PSD: its SVD is its eigen-decomposition
        lam_h = lam.cpu().numpy()
        i = 0
        while i < len(lam_h):                              # :43-48, the bound follows the scan
            self.max_energy = float(lam_h[i])
            if lam_h[i] < self.error:
                break
            i += 1
        if i == 0:
            return None
        sigma_vt = vec[:i] @ self.sketch                   # U[:, :i].T @ sketch (:52)
        vt = sigma_vt / torch.sqrt(lam[:i])[:, None]
        dv = self.sketch @ vt.T
        self.sketch = self.sketch - dv @ vt
        self.cov = self.cov - dv @ dv.T
        return sigma_vt

    def fit(self, X):
        """One row (1, d); returns the dumped rows (k x d tensor) or None (fastfdwithdump.py:91-113)."""
        x = self._row(X)
        self.max_energy += float((x * x).sum().item())
        self.sketch = torch.cat([self.sketch, x], dim=0)
        if self.sketch.shape[0] >= 2 * self.sketch_dim:
            # full FD shrink of the 2 l stacked rows (:98-108); sigma^2 comes from the same decomposition
            lib = _lib.load()
            out = torch.empty((self.sketch_dim, self.d), dtype=torch.float64, device=self._dev)
            s2 = torch.empty(self.sketch_dim, dtype=torch.float64, device=self._dev)
            stack = self.sketch.contiguous()
            with torch.cuda.device(self._dev):
                _lib.check(lib.mused_fd_merge(stack.data_ptr(), stack.shape[0], self.d, self.d, self.sketch_dim,
                                              out.data_ptr(), s2.data_ptr(), None, torch.cuda.current_stream().cuda_stream))
            self.sketch, self.sigma_squared = out, s2
            return self.dump()
        return self._border_and_deflate()

    def dump(self):
        """fastfdwithdump.py:116-133: every row with sigma^2 >= threshold leaves; the Gram is rebuilt."""
        idx = torch.nonzero(self.sigma_squared >= self.error).flatten()
        ret = None
        if idx.numel() > 0:
            ret = self.sketch[idx].clone()
            self.sketch[idx] = 0.0
            self.sigma_squared[idx] = 0.0
            self.max_energy = float(self.sigma_squared.max().item())
        self.cov = self.sketch @ self.sketch.T
        return ret

    def get(self):
        return self.sketch.cpu().numpy(), None, None, None

    def get_sketch(self):
        return self.sketch


class _Snapshot:
    __slots__ = ("v", "s", "t")

    def __init__(self, v, s, t):
        self.v, self.s, self.t = v, s, t


class _Side:
    """DumpFDwithSnapshotQueue (sliding_windows_frequent_directions.py:19-24)."""

    def __init__(self, C_, last_dump_time=np.uint64(1)):
        self.C = C_
        self.queue = deque()
        self.last_dump_time = last_dump_time


class FastSlidingWindowFD:
    def __init__(self, N, d, sketch_dim, C=1, *, device=0, **kwargs):
        self.error = kwargs["error_threshold"] if "error_threshold" in kwargs else N * 1.0 / sketch_dim
        self._device = int(device)
        self.N, self.d, self.C, self.sketch_dim = np.uint64(N), int(d), C, int(sketch_dim)
        self.fd = _Side(self._fresh())
        self.fd_aux = _Side(self._fresh())
        self.period = 0
        self.time = np.uint64(0)

    def _fresh(self):
        return FastFrequentDirectionsWithDump(self.d, self.sketch_dim * self.C, self.error, device=self._device)

    def _tick(self, t):
        self.time = np.uint64(t) if t is not None else self.time + np.uint64(1)
        while self.period < self.time // self.N:           # period swap (:233-241, :265-273)
            self.fd = self.fd_aux
            self.fd_aux = _Side(self._fresh(), last_dump_time=self.time)
            self.period += 1

    def _push(self, side, v):
        s = side.last_dump_time
        side.last_dump_time = self.time + 1
        side.queue.append(_Snapshot(v, s, self.time))

    def append(self, X, t=None):
        """Raw-row snapshot on both sides (:223-251)."""
        self._tick(t)
        x = self.fd.C._row(X)
        self._push(self.fd, x)
        self._push(self.fd_aux, x)

    def fit(self, X, t=None):
        """(:254-303)"""
        self._tick(t)
        for side in (self.fd, self.fd_aux):
            while side.queue and side.queue[0].t + self.N <= self.time:
                side.queue.popleft()
        for side in (self.fd, self.fd_aux):
            dumped = side.C.fit(X)
            if dumped is not None:
                self._push(side, dumped)

    def get(self):
        """(:306-327) the current stack under every live snapshot's rows, FD-shrunk to l rows."""
        q = self.fd.queue
        Cs = self.fd.C.get_sketch()
        if len(q) != 0:
            stack = torch.cat([Cs] + [s.v for s in q], dim=0)
            out = fd_merge(stack, self.sketch_dim)
            return out.cpu().numpy(), None, None, 0.0
        return self.fd.C.get()

    def get_size(self):
        return len(self.fd.queue) + len(self.fd_aux.queue) + 2 * self.sketch_dim


class FastTimeBasedSWFD:
    def __init__(self, N, R, d, sketch_dim, beta=4.0, *, device=0, **kwargs):
        self.N, self.d, self.R, self.beta, self.sketch_dim = np.uint64(N), int(d), R, beta, int(sketch_dim)
        self.log_eNR = int(np.ceil(np.log2(N * R / sketch_dim)))                   # seq_based_swfd.py:222
        base = 1.0
        if "upper_F_norm" in kwargs:                                               # :233-240
            self.log_eNR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / self.sketch_dim)))
            if "lower_F_norm" in kwargs:
                self.log_eNR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / kwargs["lower_F_norm"])))
                base = kwargs["lower_F_norm"] / self.sketch_dim
        self.levels = {j: FastSlidingWindowFD(self.N, self.d, self.sketch_dim, device=device, error_threshold=base * (2 ** j))
                       for j in range(self.log_eNR)}
        self.time = np.uint64(0)

    def fit(self, X, t):
        """Row X arriving at time t (:253-289)."""
        self.time = t
        x = np.asarray(X.cpu() if isinstance(X, torch.Tensor) else X, dtype=np.float64).reshape(1, self.d)
        e = float((x @ x.T).item())
        cap = (2 + 8 / self.beta) * self.sketch_dim
        for j in range(self.log_eNR):
            lv = self.levels[j]
            for side in (lv.fd, lv.fd_aux):
                q = side.queue
                while q and (len(q) > cap or q[0].t + self.N <= self.time):
                    q.popleft()
            if e >= lv.fd_aux.C.get_error():
                lv.append(x, t)
            else:
                lv.fit(x, t)

    def get(self):
        """(:291-308)"""
        rj = 0
        for j in range(self.log_eNR):
            q = self.levels[j].fd.queue
            if len(q) == 0:
                break
            if self.time - q[0].s >= min(self.N - 1, self.time - 1):
                rj = j
                break
        return self.levels[rj].get()

    def get_size(self):
        return sum(self.levels[j].get_size() for j in self.levels)

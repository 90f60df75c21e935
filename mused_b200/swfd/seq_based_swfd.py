"""SeqBasedSWFD on the B200 engine -- same constructor, ``fit``, ``get`` and
``get_size`` as /root/reference/swfd/seq_based_swfd.py:11-103.

Host code only sizes the level ladder (``logR``, thresholds) exactly like the
reference constructor and forwards rows; every per-row decision and all
arithmetic run in the CUDA engine (csrc/swfd_device.cuh) through the C ABI.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib


class SeqBasedSWFD:
    def __init__(self, N: int, R: float, d: int, sketch_dim: int, beta: float = 1.0, *,
                 device: int = 0, cluster: int = 0, chunk_rows: int = 0,
                 schedule: str = "reference", batch_rows: int = 0, n_streams: int = 1, **kwargs):
        """Sliding Window on Frequent Directions (seq_based_swfd.py:12-44).

        Args (reference): N window size, R upper bound of squared row norm,
        d row dimension, sketch_dim, beta; kwarg ``upper_F_norm`` switches the
        level count to ceil(log2(F/N)) (:31-37).
        Extras (keyword-only): device ordinal, cluster (CTAs per FD instance,
        0 = auto), chunk_rows (rows staged per launch), schedule:
          "reference"  every flush / dump decision of the reference, row by row
                       (float64, bit-faithful schedule; the parity path);
          "batched"    one tensor-core Gram + eigensolve + rotate per chunk of
                       batch_rows (default sketch_dim) rows, all directions with
                       sigma^2 >= threshold dumped per chunk (the throughput path;
                       semantics stated in oracle/swfd_batched_oracle.py); sketch_dim
                       2..256, any d > sketch_dim.
        """
        self.N = np.uint64(N)
        self.d = int(d)
        self.R = R
        self.logR = int(np.ceil(np.log2(R)))                       # :25
        self.beta = beta
        self.sketch_dim = int(sketch_dim)
        if "upper_F_norm" in kwargs:                               # :31-37
            self.logR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / self.N)))
            base = self.N / self.sketch_dim
            thr = [float(base * (2 ** j)) for j in range(self.logR)]
        else:                                                      # :39-42
            thr = [float((2 ** j) * self.N / self.sketch_dim) for j in range(self.logR)]
        self.logR = max(self.logR, 0) if self.logR < 0 else self.logR
        self.thresholds = thr
        lib = _lib.load()
        arr = (C.c_double * max(1, len(thr)))(*thr)
        if schedule not in _lib.SCHEDULES:
            raise ValueError("schedule must be 'reference' or 'batched', got %r" % (schedule,))
        self.schedule = schedule
        cfg = _lib.SwfdCfg(int(N), self.d, self.sketch_dim, float(beta), len(thr), arr,
                           int(device), int(cluster), int(chunk_rows),
                           _lib.SCHEDULES[schedule], int(batch_rows), int(n_streams))
        self.n_streams = int(n_streams)
        self._h = C.c_void_p()
        _lib.check(lib.mused_swfd_create(C.byref(self._h), C.byref(cfg)))
        self._lib = lib
        self._device = int(device)
        self._cfg_key = dict(N=int(N), d=self.d, sketch_dim=self.sketch_dim, beta=float(beta), thresholds=list(thr),
                             schedule=schedule, batch_rows=int(batch_rows), n_streams=int(n_streams),
                             cluster=int(cluster), chunk_rows=int(chunk_rows))

    # -- checkpoint / resume (no counterpart in the reference: SURVEY.md section 5 "state_dict") ---------------
    def state_dict(self):
        """The sketch as a picklable dict: the configuration it was built with and one uint8 array holding the
        engine's device state (C ABI ``mused_swfd_state_save``).  ``load_state_dict`` on a sketch constructed with
        the same arguments continues the stream exactly where this one stands (bit-identical results)."""
        n = C.c_int64(0)
        _lib.check(self._lib.mused_swfd_state_size(self._h, C.byref(n)))
        blob = np.empty(int(n.value), dtype=np.uint8)
        _lib.check(self._lib.mused_swfd_state_save(self._h, blob.ctypes.data, int(n.value)))
        return dict(format="mused_b200.swfd/1", cfg=dict(self._cfg_key), blob=blob,
                    last_level=getattr(self, "last_level", None))

    def load_state_dict(self, state):
        if state.get("format") != "mused_b200.swfd/1":
            raise ValueError("not a mused_b200 sketch state")
        if state["cfg"] != self._cfg_key:
            diff = {k: (state["cfg"].get(k), v) for k, v in self._cfg_key.items() if state["cfg"].get(k) != v}
            raise ValueError("state was saved from a sketch with another configuration: %r" % (diff,))
        blob = np.ascontiguousarray(np.asarray(state["blob"], dtype=np.uint8))
        _lib.check(self._lib.mused_swfd_state_load(self._h, blob.ctypes.data, int(blob.size)))
        if state.get("last_level") is not None:
            self.last_level = state["last_level"]

    # -- reference API -----------------------------------------------------
    def fit(self, X):
        """Handle the arriving row(s) (seq_based_swfd.py:47-80).

        X: (1, d) like the reference, or (n, d) = n consecutive arrivals in
        order (identical to n single calls).  numpy (host) or torch (host or
        CUDA) float64 / float32 / int64 / int32 / uint8 rows; copied.
        """
        keep, ptr, n, ld, dt, where = _lib.rows_arg(X, self.d)
        _lib.check(self._lib.mused_swfd_fit(self._h, ptr, n, ld, dt, where,
                                            _lib.current_stream(getattr(self, "_device", 0))))
        del keep

    def get(self):
        """(sketch l x d, sigma_squared l, Vt l x d, delta) -- seq_based_swfd.py:83-100."""
        ell, d = self.sketch_dim, self.d
        sketch = np.empty((ell, d))
        s2 = np.empty(ell)
        vt = np.empty((ell, d))
        delta = C.c_double(0.0)
        level = C.c_int32(0)
        if self.logR <= 0:
            raise KeyError(0)                                      # what the reference does
        _lib.check(self._lib.mused_swfd_get(self._h, sketch.ctypes.data, s2.ctypes.data,
                                            vt.ctypes.data, C.byref(delta), C.byref(level),
                                            _lib.HOST, _lib.current_stream(getattr(self, "_device", 0))))
        self.last_level = int(level.value)
        return sketch, s2, vt, float(delta.value)

    def get_size(self):
        n = C.c_int64(0)
        _lib.check(self._lib.mused_swfd_size(self._h, C.byref(n)))
        return int(n.value)

    # -- extras --------------------------------------------------------------
    @property
    def time(self):
        t = C.c_uint64(0)
        _lib.check(self._lib.mused_swfd_time(self._h, C.byref(t)))
        return np.uint64(t.value)

    def debug_state(self, stream=None):
        """Per FD instance state in the layout of oracle.swfd_oracle.SeqSWFD.debug_state
        (``stream`` = index of the stream of a MultiStreamSWFD)."""
        S = 2 * len(self.thresholds)
        cap = C.c_int64(0)
        _lib.check(self._lib.mused_swfd_queue_capacity(self._h, C.byref(cap)))
        cap = int(cap.value)
        scal = np.zeros((max(S, 1), 8))
        s2 = np.zeros((max(S, 1), self.sketch_dim))
        qs = np.zeros((max(S, 1), cap), dtype=np.uint64)
        qt = np.zeros((max(S, 1), cap), dtype=np.uint64)
        if stream is None:
            _lib.check(self._lib.mused_swfd_debug(self._h, scal.ctypes.data, s2.ctypes.data,
                                                  qs.ctypes.data, qt.ctypes.data, cap))
        else:
            _lib.check(self._lib.mused_swfd_debug_stream(self._h, int(stream), scal.ctypes.data, s2.ctypes.data,
                                                         qs.ctypes.data, qt.ctypes.data, cap))
        out = []
        for o in range(S):
            ql = int(scal[o, 4])
            out.append(dict(level=o // 2, energy=float(scal[o, 0]), delta=float(scal[o, 1]),
                            pending=int(scal[o, 2]), last_dump_time=int(scal[o, 3]),
                            s2=s2[o].copy(), q_s=[int(v) for v in qs[o, :ql]],
                            q_t=[int(v) for v in qt[o, :ql]], n_flush=int(scal[o, 5]),
                            n_dump=int(scal[o, 6]), n_raw=int(scal[o, 7])))
        return out

    PHASES = ("A_book", "B_gemv", "C_QTg", "D_deflate", "E_roots", "F_what", "G_urows", "H_gemm",
              "flush", "dump", "raw", "rows")

    def profile(self):
        """SM cycles per phase of the per-row kernel, per physical FD instance."""
        S = max(1, 2 * len(self.thresholds))
        buf = np.zeros((S, 16), dtype=np.uint64)
        _lib.check(self._lib.mused_swfd_profile(self._h, buf.ctypes.data, S))
        return [dict(zip(self.PHASES, (int(v) for v in row[:12]))) for row in buf[:2 * len(self.thresholds)]]

    def kernel_timing(self, enable=-1):
        """(ms, launches, rows) of the per-row kernel since timing was armed; enable=1 arms."""
        ms, nl, nr = C.c_double(0.0), C.c_int64(0), C.c_int64(0)
        _lib.check(self._lib.mused_swfd_timing(self._h, int(enable), C.byref(ms), C.byref(nl), C.byref(nr)))
        return float(ms.value), int(nl.value), int(nr.value)

    KERNELS = ("K1_stage", "K5_book", "K2_gram", "K3_eig", "K4_rotate")

    def kernel_ms(self, enable=-1):
        """schedule="batched": ({kernel class: ms}, chunks, rows) since timing was armed."""
        ms = (C.c_double * 5)()
        nc, nr = C.c_int64(0), C.c_int64(0)
        _lib.check(self._lib.mused_swfd_kernel_ms(self._h, int(enable), ms, C.byref(nc), C.byref(nr)))
        return dict(zip(self.KERNELS, (float(v) for v in ms))), int(nc.value), int(nr.value)

    def kernel_gaps(self):
        """schedule="batched": stream idle ms (inside chunks, between chunks) of the period closed
        by the last ``kernel_ms(enable >= 0)``."""
        g = (C.c_double * 2)()
        _lib.check(self._lib.mused_swfd_kernel_gaps(self._h, g))
        return float(g[0]), float(g[1])

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.mused_swfd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiStreamSWFD(SeqBasedSWFD):
    """K independent SeqBasedSWFD sketches behind one handle (north_star: "independent streams and windows
    are batched per GPU"; BASELINE configs[4] "64 independent windows batched per GPU").  Every stream has the
    semantics of ``SeqBasedSWFD(N, R, d, sketch_dim, beta, schedule="batched")`` on its own rows
    (seq_based_swfd.py:11-103) and shares nothing with the others; the engine advances all of them in
    lock-step, one K1..K5 launch sequence per chunk for the K x 2 x levels FD instances
    (C ABI: mused_swfd_cfg.n_streams, mused_swfd_fit_streams / get_stream / size_stream)."""

    def __init__(self, n_streams: int, N: int, R: float, d: int, sketch_dim: int, beta: float = 1.0, *,
                 device: int = 0, batch_rows: int = 0, **kwargs):
        super().__init__(N, R, d, sketch_dim, beta, device=device, schedule="batched", batch_rows=batch_rows,
                         n_streams=n_streams, **kwargs)

    def fit(self, X):
        """X: (K, n, d) -- n consecutive rows for every stream (numpy or torch, host or CUDA)."""
        try:
            import torch
        except Exception:  # pragma: no cover
            torch = None
        K = self.n_streams
        if torch is not None and isinstance(X, torch.Tensor):
            if X.dim() != 3 or X.shape[0] != K or X.shape[2] != self.d:
                raise ValueError("expected (n_streams=%d, n, d=%d) rows, got %s" % (K, self.d, tuple(X.shape)))
            if X.stride(2) != 1 or (X.shape[1] > 1 and X.stride(1) < self.d):
                X = X.contiguous()
            keep, ptr, n, ld, dt, where = _lib.rows_arg(X[0], self.d)
            stride = X.stride(0) if K > 1 else n * ld
            keep = X
        else:
            X = np.asarray(X)
            if X.ndim != 3 or X.shape[0] != K or X.shape[2] != self.d:
                raise ValueError("expected (n_streams=%d, n, d=%d) rows, got %s" % (K, self.d, X.shape))
            if X.dtype not in _lib._NP_DTYPES:
                X = X.astype(np.float64)
            X = np.ascontiguousarray(X)
            keep, ptr, n, ld, dt, where = _lib.rows_arg(X[0], self.d)
            stride = X.strides[0] // X.itemsize
            keep = X
        _lib.check(self._lib.mused_swfd_fit_streams(self._h, ptr, n, ld, stride, dt, where,
                                                    _lib.current_stream(self._device)))
        del keep

    def get(self, stream=0):
        """(sketch, sigma_squared, Vt, delta) of stream ``stream`` (seq_based_swfd.py:83-100)."""
        ell, d = self.sketch_dim, self.d
        sketch, s2, vt = np.empty((ell, d)), np.empty(ell), np.empty((ell, d))
        delta, level = C.c_double(0.0), C.c_int32(0)
        if self.logR <= 0:
            raise KeyError(0)
        _lib.check(self._lib.mused_swfd_get_stream(self._h, int(stream), sketch.ctypes.data, s2.ctypes.data,
                                                   vt.ctypes.data, C.byref(delta), C.byref(level), _lib.HOST,
                                                   _lib.current_stream(self._device)))
        self.last_level = int(level.value)
        return sketch, s2, vt, float(delta.value)

    def get_size(self, stream=0):
        n = C.c_int64(0)
        _lib.check(self._lib.mused_swfd_size_stream(self._h, int(stream), C.byref(n)))
        return int(n.value)


class TimeBasedSWFD(SeqBasedSWFD):
    """TimeBasedSWFD on the B200 engine (swfd/seq_based_swfd.py:106-207): rows carry an explicit
    arrival time, the ladder has ceil(log2(N R / sketch_dim)) levels with dump thresholds 2^j
    (kwargs ``upper_F_norm`` / ``lower_F_norm`` as in :131-141), beta defaults to 4."""

    def __init__(self, N: int, R: float, d: int, sketch_dim: int, beta: float = 4.0, *,
                 device: int = 0, cluster: int = 0, chunk_rows: int = 0, **kwargs):
        self.N = np.uint64(N)
        self.d = int(d)
        self.R = R
        self.beta = beta
        self.sketch_dim = int(sketch_dim)
        self.log_eNR = int(np.ceil(np.log2(N * R / sketch_dim)))                       # :120
        base = 1.0
        if "upper_F_norm" in kwargs:                                                   # :131-141
            self.log_eNR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / self.sketch_dim)))
            if "lower_F_norm" in kwargs:
                self.log_eNR = int(np.ceil(np.log2(kwargs["upper_F_norm"] / kwargs["lower_F_norm"])))
                base = kwargs["lower_F_norm"] / self.sketch_dim
        thr = [float(base * (2 ** j)) for j in range(self.log_eNR)]                    # :139-146
        self.logR = self.log_eNR
        self.thresholds = thr
        self.schedule = "reference"
        lib = _lib.load()
        arr = (C.c_double * max(1, len(thr)))(*thr)
        cfg = _lib.SwfdCfg(int(N), self.d, self.sketch_dim, float(beta), len(thr), arr,
                           int(device), int(cluster), int(chunk_rows), _lib.SCHEDULES["reference"], 0, 1)
        self.n_streams = 1
        self._h = C.c_void_p()
        _lib.check(lib.mused_swfd_create(C.byref(self._h), C.byref(cfg)))
        self._lib = lib
        self._device = int(device)

    def fit(self, X, t):
        """Row(s) X arriving at time(s) t (:150-189): t is one integer for a (1, d) row, or a
        non-decreasing sequence of n integers for (n, d) rows (= n single calls)."""
        keep, ptr, n, ld, dt, where = _lib.rows_arg(X, self.d)
        times = np.ascontiguousarray(np.broadcast_to(np.asarray(t, dtype=np.uint64).reshape(-1), (n,))
                                     if np.ndim(t) == 0 else np.asarray(t, dtype=np.uint64).reshape(-1))
        if times.shape[0] != n:
            raise ValueError("%d arrival times for %d rows" % (times.shape[0], n))
        _lib.check(self._lib.mused_swfd_fit_timed(self._h, ptr, n, ld, dt, where, times.ctypes.data,
                                                  _lib.current_stream(getattr(self, "_device", 0))))
        del keep

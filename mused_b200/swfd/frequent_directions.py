"""RobustFrequentDirections / FrequentDirectionsWithDump on the B200 engine --
API of /root/reference/swfd/frequent_directions.py:79-153."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from .. import _lib


class _FDBase:
    def __init__(self, d, sketch_dim, error, device=0, cluster=0):
        self.d = int(d)
        self.sketch_dim = int(sketch_dim)
        self._lib = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self._lib.mused_fd_create(C.byref(self._h), self.d, self.sketch_dim,
                                             float(error), int(device), int(cluster)))

    def _get(self):
        ell, d = self.sketch_dim, self.d
        sketch = np.empty((ell, d))
        s2 = np.empty(ell)
        vt = np.empty((ell, d))
        delta = C.c_double(0.0)
        _lib.check(self._lib.mused_fd_get(self._h, sketch.ctypes.data, s2.ctypes.data,
                                          vt.ctypes.data, C.byref(delta), _lib.HOST, None))
        return sketch, s2, vt, float(delta.value)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.mused_fd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RobustFrequentDirections(_FDBase):
    """frequent_directions.py:79-93 with the base ``fit(X, batch_size)`` of :39-58."""

    def __init__(self, d, sketch_dim=8, *, device=0, cluster=0):
        super().__init__(d, sketch_dim, math.inf, device, cluster)

    def fit(self, X, batch_size=1):
        keep, ptr, n, ld, dt, where = _lib.rows_arg(X, self.d)
        _lib.check(self._lib.mused_fd_fit_batches(self._h, ptr, n, ld, dt, where, int(batch_size), None))
        del keep

    def get(self):
        return self._get()

    def get_sketch(self):
        return self._get()[0]


class FrequentDirectionsWithDump(_FDBase):
    """frequent_directions.py:96-153."""

    def __init__(self, d: int, sketch_dim: int, error: float, *, device=0, cluster=0):
        super().__init__(d, min(int(sketch_dim), int(d) - 0), error, device, cluster)
        self.error = float(error)

    def get_error(self) -> float:
        return self.error

    def fit(self, X, batch_size=1):
        keep, ptr, n, ld, dt, where = _lib.rows_arg(X, self.d)
        _lib.check(self._lib.mused_fd_fit(self._h, ptr, n, ld, dt, where, None))
        del keep

    def dump(self):
        v = np.empty((1, self.d))
        flag = C.c_int32(0)
        _lib.check(self._lib.mused_fd_dump(self._h, v.ctypes.data, C.byref(flag), _lib.HOST, None))
        return v if flag.value else None

    def flush(self):
        _lib.check(self._lib.mused_fd_flush(self._h, None))

    def get(self):
        return self._get()

    def get_sketch(self):
        return self._get()[0]

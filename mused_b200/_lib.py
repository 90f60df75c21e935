"""ctypes binding of libmused_b200.so (include/mused_b200.h).

There is no CPU fallback: if the library is missing ``load()`` raises, and on
a machine without a CUDA device every handle constructor raises RuntimeError
with the library's own message.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmused_b200.so")

F64, F32, I64, I32, U8 = 0, 1, 2, 3, 4
HOST, DEVICE = 0, 1
METRIC_COSINE, METRIC_L2, METRIC_HAVERSINE, METRIC_L1SUM = 0, 1, 2, 3

_NP_DTYPES = {np.dtype(np.float64): F64, np.dtype(np.float32): F32, np.dtype(np.int64): I64,
              np.dtype(np.int32): I32, np.dtype(np.uint8): U8, np.dtype(np.bool_): U8}


class SwfdCfg(C.Structure):
    _fields_ = [("N", C.c_int64), ("d", C.c_int64), ("ell", C.c_int64), ("beta", C.c_double),
                ("n_levels", C.c_int32), ("thresholds", C.POINTER(C.c_double)),
                ("device", C.c_int32), ("cluster", C.c_int32), ("chunk_rows", C.c_int64),
                ("schedule", C.c_int32), ("batch_rows", C.c_int32), ("n_streams", C.c_int32)]


SCHEDULES = {"reference": 0, "batched": 1}


_lib = None

# every symbol include/mused_b200.h declares
SYMBOLS = [
    "mused_last_error", "mused_version", "mused_device_count", "mused_launch_count",
    "mused_count_launches", "mused_swfd_profile", "mused_swfd_timing", "mused_swfd_kernel_ms",
    "mused_swfd_create", "mused_swfd_destroy", "mused_swfd_fit", "mused_swfd_get",
    "mused_swfd_size", "mused_swfd_time", "mused_swfd_debug", "mused_swfd_queue_capacity",
    "mused_swfd_chunk_room", "mused_swfd_stage", "mused_swfd_chunk_gram", "mused_swfd_chunk_finish",
    "mused_swfd_exchange_buffers", "mused_swfd_get_begin", "mused_swfd_get_finish",
    "mused_fd_create", "mused_fd_destroy", "mused_fd_fit", "mused_fd_fit_batches",
    "mused_fd_dump", "mused_fd_flush", "mused_fd_get",
    "mused_assign", "mused_topk_adjacency", "mused_contingency", "mused_sym_eig",
    "mused_topk_adjacency_f64", "mused_swfd_kernel_gaps", "mused_swfd_fit_timed",
    "mused_kmeans", "mused_fd_merge", "mused_comm_unique_id", "mused_swfd_attach_comm", "mused_swfd_comm_stats",
    "mused_swfd_fit_streams", "mused_swfd_get_stream", "mused_swfd_size_stream", "mused_swfd_debug_stream",
    "mused_topk_adjacency_static", "mused_assign_pending", "mused_svd_reduce",
    "mused_swfd_state_size", "mused_swfd_state_save", "mused_swfd_state_load",
    "mused_group_adjacency", "mused_jaccard_topk_adjacency",
]


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libmused_b200.so is not built (%s). Run `python -m mused_b200.build`; "
            "mused_b200 has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i64, i32, dbl = C.c_void_p, C.c_int64, C.c_int32, C.c_double
    lib.mused_last_error.restype = C.c_char_p
    lib.mused_launch_count.restype = C.c_uint64
    lib.mused_count_launches.argtypes = [C.c_uint64]
    lib.mused_count_launches.restype = None
    lib.mused_swfd_profile.argtypes = [vp, vp, i64]
    lib.mused_swfd_timing.argtypes = [vp, C.c_int, C.POINTER(dbl), C.POINTER(i64), C.POINTER(i64)]
    lib.mused_swfd_kernel_ms.argtypes = [vp, C.c_int, C.POINTER(dbl), C.POINTER(i64), C.POINTER(i64)]
    lib.mused_swfd_kernel_gaps.argtypes = [vp, C.POINTER(dbl)]
    lib.mused_swfd_chunk_room.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    lib.mused_swfd_stage.argtypes = [vp, vp, i64, i64, C.c_int, C.c_int, vp]
    lib.mused_swfd_chunk_gram.argtypes = [vp, vp]
    lib.mused_swfd_chunk_finish.argtypes = [vp, vp]
    lib.mused_swfd_exchange_buffers.argtypes = [vp, C.POINTER(vp), C.POINTER(i64), C.POINTER(vp), C.POINTER(i64),
                                                C.POINTER(vp), C.POINTER(i64)]
    lib.mused_swfd_get_begin.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), vp]
    lib.mused_swfd_get_finish.argtypes = [vp, vp, vp, vp, C.POINTER(dbl), C.c_int, vp]
    lib.mused_swfd_create.argtypes = [C.POINTER(vp), C.POINTER(SwfdCfg)]
    lib.mused_swfd_destroy.argtypes = [vp]
    lib.mused_swfd_fit.argtypes = [vp, vp, i64, i64, C.c_int, C.c_int, vp]
    lib.mused_swfd_fit_timed.argtypes = [vp, vp, i64, i64, C.c_int, C.c_int, vp, vp]
    lib.mused_swfd_get.argtypes = [vp, vp, vp, vp, C.POINTER(dbl), C.POINTER(i32), C.c_int, vp]
    lib.mused_swfd_fit_streams.argtypes = [vp, vp, i64, i64, i64, C.c_int, C.c_int, vp]
    lib.mused_swfd_get_stream.argtypes = [vp, i32, vp, vp, vp, C.POINTER(dbl), C.POINTER(i32), C.c_int, vp]
    lib.mused_swfd_size_stream.argtypes = [vp, i32, C.POINTER(i64)]
    lib.mused_swfd_debug_stream.argtypes = [vp, i32, vp, vp, vp, vp, i64]
    lib.mused_comm_unique_id.argtypes = [vp, i64, C.POINTER(i64), C.c_char_p]
    lib.mused_swfd_attach_comm.argtypes = [vp, vp, i64, i32, i32, C.c_char_p]
    lib.mused_swfd_comm_stats.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.mused_swfd_size.argtypes = [vp, C.POINTER(i64)]
    lib.mused_swfd_time.argtypes = [vp, C.POINTER(C.c_uint64)]
    lib.mused_swfd_debug.argtypes = [vp, vp, vp, vp, vp, i64]
    lib.mused_swfd_queue_capacity.argtypes = [vp, C.POINTER(i64)]
    lib.mused_fd_create.argtypes = [C.POINTER(vp), i64, i64, dbl, C.c_int, C.c_int]
    lib.mused_fd_destroy.argtypes = [vp]
    lib.mused_fd_fit.argtypes = [vp, vp, i64, i64, C.c_int, C.c_int, vp]
    lib.mused_fd_fit_batches.argtypes = [vp, vp, i64, i64, C.c_int, C.c_int, i64, vp]
    lib.mused_fd_dump.argtypes = [vp, vp, C.POINTER(i32), C.c_int, vp]
    lib.mused_fd_flush.argtypes = [vp, vp]
    lib.mused_fd_get.argtypes = [vp, vp, vp, vp, C.POINTER(dbl), C.c_int, vp]
    lib.mused_assign.argtypes = [vp, i64, i64, i64, vp, i64, i64, C.c_int, C.c_float, vp, vp, vp]
    lib.mused_swfd_state_size.argtypes = [vp, C.POINTER(i64)]
    lib.mused_swfd_state_save.argtypes = [vp, vp, i64]
    lib.mused_swfd_state_load.argtypes = [vp, vp, i64]
    lib.mused_group_adjacency.argtypes = [vp, i64, vp, vp]
    lib.mused_jaccard_topk_adjacency.argtypes = [vp, vp, i64, i64, vp, vp]
    lib.mused_svd_reduce.argtypes = [vp, C.c_int, i64, i64, i64, vp, i32, i32, i32, vp, i64, vp, vp, i64, vp]
    lib.mused_assign_pending.argtypes = [C.POINTER(C.c_uint64)]
    lib.mused_topk_adjacency.argtypes = [vp, i64, i64, i64, C.c_int, i64, vp, vp]
    lib.mused_topk_adjacency_static.argtypes = [vp, i64, i64, i64, C.c_int, i64, vp, vp]
    lib.mused_fd_merge.argtypes = [vp, i64, i64, i64, i64, vp, vp, vp, vp]
    lib.mused_kmeans.argtypes = [vp, i64, i64, i64, i32, C.c_uint64, i32, dbl, vp, vp, C.POINTER(dbl), C.POINTER(i32), vp]
    lib.mused_contingency.argtypes = [vp, vp, i64, i32, i32, vp, vp]
    lib.mused_topk_adjacency_f64.argtypes = [vp, i64, i64, i64, C.c_int, i64, vp, vp]
    lib.mused_sym_eig.argtypes = [vp, i64, i64, vp, vp, i64, C.POINTER(i32), vp]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().mused_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError(msg)
        raise RuntimeError("mused_b200: %s (code %d)" % (msg, rc))


def current_stream(device=0):
    """cudaStream_t (as int) of torch's current stream on ``device``; None (= the default stream)
    when torch or CUDA is not there.  Independent sketches driven from different torch streams
    overlap on the GPU (their kernels leave most SMs idle: one warp per SM sub-partition)."""
    try:
        import torch
        if torch.cuda.is_available():
            return int(torch.cuda.current_stream(device).cuda_stream) or None
    except Exception:  # pragma: no cover
        pass
    return None


def rows_arg(X, d):
    """(pointer, n_rows, ld, dtype, where) for a 2-D numpy array or torch tensor."""
    try:
        import torch
    except Exception:  # pragma: no cover
        torch = None
    if torch is not None and isinstance(X, torch.Tensor):
        if X.dim() == 1:
            X = X.reshape(1, -1)
        if X.shape[1] != d:
            raise ValueError("row dimension %d != d=%d" % (X.shape[1], d))
        if X.stride(1) != 1:
            X = X.contiguous()
        dt = {torch.float64: F64, torch.float32: F32, torch.int64: I64, torch.int32: I32,
              torch.uint8: U8, torch.bool: U8}.get(X.dtype)
        if dt is None:
            X = X.to(torch.float64)
            dt = F64
        where = DEVICE if X.is_cuda else HOST
        return X, X.data_ptr(), X.shape[0], X.stride(0) if X.shape[0] > 1 else d, dt, where
    A = np.asarray(X)
    if A.ndim == 1:
        A = A.reshape(1, -1)
    if A.ndim != 2 or A.shape[1] != d:
        raise ValueError("expected rows of dimension %d, got shape %s" % (d, A.shape))
    if A.dtype not in _NP_DTYPES:
        A = A.astype(np.float64)
    if A.strides[1] != A.itemsize or (A.shape[0] > 1 and A.strides[0] % A.itemsize) or not A.flags.aligned:
        A = np.ascontiguousarray(A)
    ld = A.strides[0] // A.itemsize if A.shape[0] > 1 else d
    if ld < d:
        A = np.ascontiguousarray(A)
        ld = d
    return A, A.ctypes.data, A.shape[0], ld, _NP_DTYPES[A.dtype], HOST

"""Column-sharded SeqBasedSWFD over ``torch.distributed``.

Rank r of the process group holds columns ``[lo_r, hi_r)`` of every incoming row,
of every FD sketch and of every snapshot.  The reference has no counterpart (it
is single-process); the API is the reference's ``SeqBasedSWFD`` surface
(/root/reference/swfd/seq_based_swfd.py:11-103) and the arithmetic is the batched
schedule of ``mused_b200.swfd.SeqBasedSWFD(schedule="batched")``.

Exchanges per chunk (both sums, both on device buffers the C library exposes,
NCCL over NVLink when the group's backend is nccl):

  1. ``norm2[l]``           float64 -- the global ||x||^2 of the chunk's rows decides
                            raw-snapshot routing (seq_based_swfd.py:76-77);
  2. ``gram[S][2l][2l]``    float64 -- G = sum_r R_r R_r^T, the one real exchange per
                            shrink (the eigensolve is replicated, the rotate local).

and per query one sum of the (l+q)^2 stack Gram.  Period swaps, expiry, dumps and the
level choice depend on replicated scalars only, so the ranks stay in lock-step
without further traffic.  ``get()`` all-gathers the l x d sketch (main.py:73-76 needs
it whole on the host) unless ``gather=False``.

The driver below is backend-agnostic: it talks to a *shard engine* with the phase
protocol (``room / stage / chunk_gram / chunk_finish / get_begin / get_finish`` plus
the three exchange tensors).  The product engine is :class:`CudaShardEngine`
(C ABI, no CPU fallback); tests inject a NumPy engine to exercise this host logic
with the gloo backend on CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def shard_bounds(d: int, world: int, rank: int, align: int = 32):
    """Columns [lo, hi) of ``rank``: equal shards, multiples of ``align`` columns."""
    if d % (align * world) != 0:
        raise ValueError("d=%d must be a multiple of %d x world_size=%d for column sharding" % (d, align, world))
    w = d // world
    return rank * w, (rank + 1) * w


class _DevBuf:
    """Zero-copy view of a device buffer of the C library for torch.as_tensor."""

    def __init__(self, ptr, count, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


class CudaShardEngine:
    """One rank's shard on the B200 engine (schedule="batched") through the phase C ABI."""

    def __init__(self, N, R, d_local, sketch_dim, beta=1.0, *, device=0, **kw):
        import torch
        from .swfd import SeqBasedSWFD
        self.sk = SeqBasedSWFD(N, R, d_local, sketch_dim, beta, device=device, schedule="batched", **kw)
        self.lib, self.h = self.sk._lib, self.sk._h
        self.ell, self.d = int(sketch_dim), int(d_local)
        p = [C.c_void_p() for _ in range(3)]
        n = [C.c_int64() for _ in range(3)]
        _lib.check(self.lib.mused_swfd_exchange_buffers(self.h, C.byref(p[0]), C.byref(n[0]), C.byref(p[1]),
                                                        C.byref(n[1]), C.byref(p[2]), C.byref(n[2])))
        dev = torch.device("cuda", device)
        self.norm2, self.gram, self.query_gram = (
            torch.as_tensor(_DevBuf(p[i].value, n[i].value), device=dev) for i in range(3))

    def room(self):
        r, q = C.c_int64(), C.c_int64()
        _lib.check(self.lib.mused_swfd_chunk_room(self.h, C.byref(r), C.byref(q)))
        return int(r.value)

    def pending(self):
        r, q = C.c_int64(), C.c_int64()
        _lib.check(self.lib.mused_swfd_chunk_room(self.h, C.byref(r), C.byref(q)))
        return int(q.value)

    def stage(self, X):
        keep, ptr, n, ld, dt, where = _lib.rows_arg(X, self.d)
        _lib.check(self.lib.mused_swfd_stage(self.h, ptr, n, ld, dt, where, None))
        del keep

    def chunk_gram(self):
        _lib.check(self.lib.mused_swfd_chunk_gram(self.h, None))

    def chunk_finish(self):
        _lib.check(self.lib.mused_swfd_chunk_finish(self.h, None))

    def get_begin(self):
        lv, ng = C.c_int32(), C.c_int32()
        _lib.check(self.lib.mused_swfd_get_begin(self.h, C.byref(lv), C.byref(ng), None))
        return int(lv.value), int(ng.value)

    def get_finish(self):
        sketch, s2, vt = np.empty((self.ell, self.d)), np.empty(self.ell), np.empty((self.ell, self.d))
        delta = C.c_double(0.0)
        _lib.check(self.lib.mused_swfd_get_finish(self.h, sketch.ctypes.data, s2.ctypes.data, vt.ctypes.data,
                                                  C.byref(delta), _lib.HOST, None))
        return sketch, s2, vt, float(delta.value)

    def get_size(self):
        return self.sk.get_size()

    def debug_state(self):
        return self.sk.debug_state()

    def close(self):
        self.sk.close()


class ColumnShardedSWFD:
    """``SeqBasedSWFD(N, R, d, sketch_dim, beta)`` with d sharded over ``process_group``."""

    def __init__(self, N, R, d, sketch_dim, beta=1.0, *, process_group=None, device=None,
                 engine_factory=None, **kwargs):
        import torch.distributed as dist
        self.dist = dist
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self.N, self.d, self.sketch_dim = int(N), int(d), int(sketch_dim)
        self.lo, self.hi = shard_bounds(self.d, self.world, self.rank)
        self.logR = int(np.ceil(np.log2(R)))                         # seq_based_swfd.py:25
        if engine_factory is None:
            if device is None:
                import torch
                device = torch.cuda.current_device()
            self.engine = CudaShardEngine(N, R, self.hi - self.lo, sketch_dim, beta, device=device, **kwargs)
        else:
            self.engine = engine_factory(N, R, self.hi - self.lo, sketch_dim, beta, **kwargs)
        self.time = 0
        self.last_level = 0
        self.n_allreduce = 0

    # -- exchanges ---------------------------------------------------------------
    def _sum(self, t):
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
            self.n_allreduce += 1

    def _close_chunk(self):
        e = self.engine
        if e.pending() == 0:
            return
        self._sum(e.norm2)            # exchange 1: global ||x||^2 of the chunk rows
        e.chunk_gram()
        self._sum(e.gram)             # exchange 2: G = sum over column shards
        e.chunk_finish()

    # -- reference API -----------------------------------------------------------
    def _local(self, X):
        if hasattr(X, "dim"):
            if X.dim() == 1:
                X = X.reshape(1, -1)
        else:
            X = np.asarray(X)
            if X.ndim == 1:
                X = X.reshape(1, -1)
        if X.shape[1] == self.d and self.world > 1:
            return X[:, self.lo:self.hi]
        if X.shape[1] != self.hi - self.lo:
            raise ValueError("rows must have d=%d (global) or %d (this rank's shard) columns, got %d"
                             % (self.d, self.hi - self.lo, X.shape[1]))
        return X

    def fit(self, X):
        """Rows in arrival order, either whole (n, d) -- the shard is sliced here -- or
        already sharded (n, hi - lo).  Every rank must be fed the same rows."""
        X = self._local(X)
        n, pos = X.shape[0], 0
        while pos < n:
            m = min(self.engine.room(), n - pos)
            self.engine.stage(X[pos:pos + m])
            pos += m
            if self.engine.room() == 0:
                self._close_chunk()
        self.time += n

    def get(self, gather=True):
        if self.logR <= 0:
            raise KeyError(0)
        self._close_chunk()           # get() closes the open chunk (the reference flushes on get)
        level, n_stack = self.engine.get_begin()
        self.last_level = level
        if n_stack > 0:
            self._sum(self.engine.query_gram)
        sketch, s2, vt, delta = self.engine.get_finish()
        if gather and self.world > 1:
            sketch = self._gather_cols(sketch)
            vt = self._gather_cols(vt)
        return sketch, s2, vt, delta

    def _gather_cols(self, A):
        import torch
        t = torch.as_tensor(np.ascontiguousarray(A))
        dev = getattr(self.engine, "norm2").device
        t = t.to(dev)
        parts = [torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(parts, t, group=self.group)
        return torch.cat(parts, dim=1).cpu().numpy()

    def get_size(self):
        return self.engine.get_size()

    def debug_state(self):
        return self.engine.debug_state()

    def close(self):
        if hasattr(self.engine, "close"):
            self.engine.close()


def find_nccl_library():
    """The libnccl.so.2 the process should share with torch (its bundled copy); None lets the library fall back
    to whatever "libnccl.so.2" resolves to."""
    import importlib.util
    import os
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except Exception:
        spec = None
    if spec is not None and spec.submodule_search_locations:
        p = os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2")
        if os.path.exists(p):
            return p
    return None


class NcclShardedSWFD:
    """``SeqBasedSWFD(N, R, d, sketch_dim, beta)`` with d column-sharded over the ranks of ``process_group`` and
    BOTH exchanges inside the C library (north_star: "an NCCL-over-NVLink allreduce of only the
    Gram per shrink"): rank 0 draws a NCCL unique id, torch.distributed broadcasts its 128 bytes, every rank
    attaches it to its engine; ``fit`` / ``get`` are then collective calls whose per-chunk allreduces
    (||x||^2, Gram) are posted on the engine's CUDA stream between its kernels -- no Python in the chunk loop,
    no host synchronisation.  ``torch.distributed`` is plumbing only (bootstrap + the optional all-gather of the
    result).  Decisions depend on replicated scalars, so the ranks stay in lock-step (seq_based_swfd.py:76-80).
    """

    def __init__(self, N, R, d, sketch_dim, beta=1.0, *, process_group=None, device=None, **kwargs):
        import torch
        import torch.distributed as dist
        from .swfd import SeqBasedSWFD
        self.dist, self.group = dist, process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self.N, self.d, self.sketch_dim = int(N), int(d), int(sketch_dim)
        self.lo, self.hi = shard_bounds(self.d, self.world, self.rank)
        if device is None:
            device = torch.cuda.current_device()
        self.device = int(device)
        self.sk = SeqBasedSWFD(N, R, self.hi - self.lo, sketch_dim, beta, device=self.device, schedule="batched", **kwargs)
        self.logR = self.sk.logR
        self.last_level = 0
        if self.world > 1:
            lib = self.sk._lib
            path = find_nccl_library()
            cpath = path.encode() if path else None
            buf = (C.c_ubyte * 256)()
            n = C.c_int64(128)
            if self.rank == 0:
                _lib.check(lib.mused_comm_unique_id(buf, 256, C.byref(n), cpath))
            t = torch.tensor(list(buf), dtype=torch.uint8, device=torch.device("cuda", self.device))
            src = dist.get_global_rank(process_group, 0) if process_group is not None else 0
            dist.broadcast(t, src=src, group=process_group)
            raw = bytes(t.cpu().tolist())
            _lib.check(lib.mused_swfd_attach_comm(self.sk._h, raw, int(n.value), self.rank, self.world, cpath))

    def _local(self, X):
        if X.shape[-1] == self.d and self.world > 1:
            return X[:, self.lo:self.hi]
        return X

    def fit(self, X):
        """Every rank feeds the same rows: whole (n, d) rows (sliced here) or its own (n, hi - lo) columns."""
        self.sk.fit(self._local(X))

    def get(self, gather=True):
        sketch, s2, vt, delta = self.sk.get()
        self.last_level = self.sk.last_level
        if gather and self.world > 1:
            import torch
            dev = torch.device("cuda", self.device)
            out = []
            for M in (sketch, vt):
                t = torch.as_tensor(np.ascontiguousarray(M)).to(dev)
                parts = [torch.empty_like(t) for _ in range(self.world)]
                self.dist.all_gather(parts, t, group=self.group)
                out.append(torch.cat(parts, dim=1).cpu().numpy())
            sketch, vt = out
        return sketch, s2, vt, delta

    def comm_stats(self):
        a, b = C.c_uint64(0), C.c_uint64(0)
        _lib.check(self.sk._lib.mused_swfd_comm_stats(self.sk._h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def get_size(self):
        return self.sk.get_size()

    def debug_state(self):
        return self.sk.debug_state()

    def close(self):
        self.sk.close()

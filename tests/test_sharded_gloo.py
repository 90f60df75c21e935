"""CPU, world_size 2, gloo: the product's column-sharded driver
(mused_b200.sharded.ColumnShardedSWFD -- chunking, exchange order, column slicing,
gather) with the NumPy shard engine standing in for the CUDA one.  The sharded run
must reproduce the unsharded batched oracle: identical schedule (snapshot stamps,
flush / dump / raw counts, level, size), sigma and covariance error to float64 Gram
accuracy, and only the two exchanges per chunk SURVEY 8(e) names."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mused_b200 import synth
from mused_b200.sharded import ColumnShardedSWFD, shard_bounds
from oracle import swfd_oracle as so
from oracle.shard_engine_oracle import ShardEngineOracle
from oracle.swfd_batched_oracle import BatchedSeqSWFD

D, ELL, N, ROWS = 128, 16, 300, 900


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = synth.lowrank_stream(ROWS, D, seed=4, scale=2.0).astype(np.float64)
        R = synth.row_norm2_max(A)
        sk = ColumnShardedSWFD(N, R, D, ELL, engine_factory=ShardEngineOracle)
        lo, hi = shard_bounds(D, world, rank)
        assert (sk.lo, sk.hi) == (lo, hi) and hi - lo == D // world
        res = []
        pos = 0
        for step in (1, 150, 149, 250, 350):         # ragged fit() calls, whole rows (sliced by the driver)
            blk = A[pos:pos + step]
            if step == 250:
                blk = blk[:, lo:hi]                  # ... and one call with pre-sharded rows
            sk.fit(blk)
            pos += step
            if pos >= N:
                sketch, s2, vt, delta = sk.get()
                assert sketch.shape == (ELL, D)      # gathered
                res.append((pos, sketch, s2, sk.last_level, sk.get_size(), sk.debug_state()))
        if rank == 0:
            out.put((res, sk.n_allreduce))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_column_sharded_driver_matches_unsharded_oracle_world2():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res, n_allreduce = out.get(timeout=240)  # before join: the payload is larger than a pipe buffer
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0

    A = synth.lowrank_stream(ROWS, D, seed=4, scale=2.0).astype(np.float64)
    R = synth.row_norm2_max(A)
    ref = BatchedSeqSWFD(N, R, D, ELL)
    pos = 0
    qi = 0
    n_queries_with_stack = 0
    for step in (1, 150, 149, 250, 350):
        ref.fit(A[pos:pos + step])
        pos += step
        if pos >= N:
            p2, sketch, s2, level, size, state = res[qi]
            qi += 1
            so_, s2o, _, _ = ref.get()
            assert p2 == pos and level == ref.pick_level() and size == ref.get_size()
            a, b = np.sqrt(np.maximum(s2o, 0)), np.sqrt(np.maximum(s2, 0))
            big = a >= 1e-3 * a.max()
            assert np.max(np.abs(a[big] - b[big]) / a[big]) <= 1e-7
            Aw = A[pos - N:pos]
            eo, es = so.cov_error(Aw, so_), so.cov_error(Aw, sketch)
            assert abs(eo - es) <= 1e-6 * eo
            for g, o in zip(state, ref.debug_state()):
                assert g["q_s"] == o["q_s"] and g["q_t"] == o["q_t"]
                assert (g["n_flush"], g["n_dump"], g["n_raw"]) == (o["n_flush"], o["n_dump"], o["n_raw"])
                assert g["last_dump_time"] == o["last_dump_time"]
            n_queries_with_stack += 1
    assert qi == len(res) == 3
    # exchanges: 2 per chunk + at most 1 per query, nothing else
    n_chunks = 0
    t = 0
    for end in (300, 550, 900):                      # chunk boundaries: every ELL rows, at multiples of N, at get()
        while t < end:
            room = min(ELL, N - (t + 1) % N if (t + 1) % N else N, end - t)
            t += room
            n_chunks += 1
    assert 2 * n_chunks <= n_allreduce <= 2 * n_chunks + 3


def test_shard_bounds():
    assert shard_bounds(1024, 8, 3) == (384, 512)
    with pytest.raises(ValueError):
        shard_bounds(100, 2, 0)

"""mused_topk_adjacency_static: the allocation-free variant of the kNN adjacency operator (scratch in a file-scope
``static __device__`` array of the library, W <= MUSED_TOPK_STATIC_MAX_W = 1024).  It has no counterpart in the
reference; it is checked, through the C ABI, against the existing entry point mused_topk_adjacency on the same
inputs: the adjacency is uint8 0/1, so the bar is bit-equality (the bar of tests/test_ops_gpu.py for this operator).

Two calls in a row with DIFFERENT shapes and metrics: the second call runs on whatever the first left in the shared
scratch, which is what a stale-contents bug would trip on.  Collected last (test_zz_*) so that a failure here hides
nothing under ``pytest -x``; passes alone in a fresh process (first use resolves the device address) and after the
other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_W = 1024
COSINE, L2 = 0, 1


def _call(fn, F, W, f, metric, k, adj):
    from mused_b200 import _lib
    with torch.cuda.device(F.device):
        return fn(F.data_ptr(), W, f, F.stride(0), metric, k, adj.data_ptr(), torch.cuda.current_stream().cuda_stream)


def _pair(lib, W, f, metric, k, seed, prefill):
    """(existing entry point, static entry point) on the same features; both OR into the same pre-filled matrix."""
    from mused_b200 import _lib
    g = torch.Generator(device="cpu").manual_seed(seed)
    F = torch.randn((W, f), generator=g, dtype=torch.float32).cuda()
    base = (torch.rand((W, W), generator=g) < prefill).to(torch.uint8).cuda()
    a, b = base.clone(), base.clone()
    _lib.check(_call(lib.mused_topk_adjacency, F, W, f, metric, k, a))
    _lib.check(_call(lib.mused_topk_adjacency_static, F, W, f, metric, k, b))
    torch.cuda.synchronize()
    return a.cpu().numpy(), b.cpu().numpy(), base.cpu().numpy()


def test_static_scratch_adjacency_matches_allocating_entry_point_twice():
    from mused_b200 import _lib
    lib = _lib.load()
    # call 1: the largest window the static scratch takes, cosine; call 2: a smaller ragged window, Euclidean, on
    # the scratch call 1 filled; call 3: back to a large one
    for W, f, metric, k, seed, prefill in ((MAX_W, 96, COSINE, 12, 1, 0.0), (333, 40, L2, 7, 2, 0.01),
                                            (1000, 64, L2, 33, 3, 0.0), (1, 8, COSINE, 1, 4, 0.0)):
        a, b, base = _pair(lib, W, f, metric, k, seed, prefill)
        assert np.array_equal(a, b), "static-scratch adjacency differs from mused_topk_adjacency at W=%d" % W
        assert np.all(b >= base)                                   # OR-accumulates like the existing entry point
        if W > k:
            assert (b.astype(np.int64) - base).sum() > 0


def test_static_scratch_makes_no_allocation_and_rejects_large_windows():
    from mused_b200 import _lib
    lib = _lib.load()
    F = torch.randn((512, 32), dtype=torch.float32, device="cuda")
    adj = torch.zeros((512, 512), dtype=torch.uint8, device="cuda")
    _lib.check(_call(lib.mused_topk_adjacency_static, F, 512, 32, COSINE, 5, adj))    # address resolved, module loaded
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _lib.check(_call(lib.mused_topk_adjacency_static, F, 512, 32, COSINE, 5, adj))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "device memory in use changed across allocation-free calls"
    # one row over the limit: a clear error, nothing launched
    W = MAX_W + 1
    Fb = torch.randn((W, 8), dtype=torch.float32, device="cuda")
    big = torch.zeros((W, W), dtype=torch.uint8, device="cuda")
    rc = _call(lib.mused_topk_adjacency_static, Fb, W, 8, COSINE, 3, big)
    assert rc != 0
    msg = lib.mused_last_error().decode()
    assert "1024" in msg and "mused_topk_adjacency_static" in msg
    torch.cuda.synchronize()
    assert int(big.sum().item()) == 0

"""One process, two GPUs: sketches and operators created on device 0 and then on device 1 give the same results
(per-device kernel attributes -- dynamic shared memory opt-ins, non-portable cluster sizes -- and per-device scratch
addresses must not be cached process-wide).  Needs 2 GPUs; skipped otherwise."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_sketches_and_operators_on_both_devices_of_one_process():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import ctypes as C
    from mused_b200 import matrix_operations as mo, synth, _lib
    from mused_b200.swfd import SeqBasedSWFD
    d, ell, N, n = 256, 64, 600, 1500
    A = synth.lowrank_stream(n, d, seed=4, scale=2.0)
    R = synth.row_norm2_max(A)
    out = {}
    for dev in (0, 1, 0):                                          # and back: nothing of device 1 sticks either
        res = {}
        for schedule in ("reference", "batched"):
            sk = SeqBasedSWFD(N, R, d, ell, schedule=schedule, device=dev)
            sk.fit(torch.as_tensor(A).to("cuda:%d" % dev))
            B, s2, vt, delta = sk.get()
            res[schedule] = (B, s2, sk.get_size(), sk.last_level)
            sk.close()
        rng = np.random.default_rng(0)
        Cn = rng.standard_normal((40, d)).astype(np.float32)
        res["labels"] = mo.cosine_assign(A, Cn, 0.1, device="cuda:%d" % dev).cpu().numpy()
        lab3, sc3 = mo.cosine_assign(A, Cn, 0.1, device="cuda:%d" % dev, return_score=True)
        res["labels3"], res["score"] = lab3.cpu().numpy(), sc3.cpu().numpy()
        res["adj"] = mo.topk_adjacency(A[:700, :48], 9, metric="cosine", device="cuda:%d" % dev).cpu().numpy()
        fused = mo.fused_adjacency([A[:500, :32]], ["dense"], k_basis=12, device="cuda:%d" % dev)
        res["svd"] = mo.perform_svd_reduction(fused, 10, 3, device="cuda:%d" % dev)
        # the allocation-free entry point resolves its scratch address per device
        lib = _lib.load()
        F = torch.as_tensor(A[:300, :40]).to("cuda:%d" % dev).contiguous()
        a1 = torch.zeros((300, 300), dtype=torch.uint8, device=F.device)
        a2 = torch.zeros_like(a1)
        with torch.cuda.device(F.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(lib.mused_topk_adjacency(F.data_ptr(), 300, 40, 40, 0, 7, a1.data_ptr(), st))
            _lib.check(lib.mused_topk_adjacency_static(F.data_ptr(), 300, 40, 40, 0, 7, a2.data_ptr(), st))
        torch.cuda.synchronize(F.device)
        assert torch.equal(a1, a2)
        res["static"] = a2.cpu().numpy()
        out.setdefault(dev, []).append(res)
    r0, r1, r0b = out[0][0], out[1][0], out[0][1]
    for other in (r1, r0b):
        for schedule in ("reference", "batched"):
            assert np.array_equal(r0[schedule][0], other[schedule][0]) and np.array_equal(r0[schedule][1], other[schedule][1])
            assert r0[schedule][2:] == other[schedule][2:]
        for key in ("labels", "labels3", "score", "adj", "static"):
            assert np.array_equal(r0[key], other[key]), key
        assert np.abs(r0["svd"] - other["svd"]).max() <= 1e-9 * np.abs(r0["svd"]).max()

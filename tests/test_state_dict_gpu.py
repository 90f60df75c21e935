"""state_dict / load_state_dict of the sketch (C ABI mused_swfd_state_size / _save / _load): a sketch restored from a
checkpoint continues the stream bit-identically to the one that never stopped -- both schedules, the batched one
with rows pending in its open chunk, many streams, a pickle round trip -- and a blob does not load into a sketch of
another configuration."""
import io
import pickle

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _same_state(a, b, stream=None):
    sa, sb = a.debug_state(stream), b.debug_state(stream)
    for x, y in zip(sa, sb):
        for key in ("energy", "delta", "pending", "last_dump_time", "q_s", "q_t", "n_flush", "n_dump", "n_raw"):
            assert x[key] == y[key], key
        assert np.array_equal(x["s2"], y["s2"])


@pytest.mark.parametrize("schedule,n1,n2", [("reference", 700, 1500), ("batched", 731, 1500), ("batched", 640, 1300)])
def test_restored_sketch_continues_bit_identically(schedule, n1, n2):
    from mused_b200 import synth
    from mused_b200.swfd import SeqBasedSWFD
    d, ell, N = 96, 16, 400
    A = synth.lowrank_stream(n2, d, seed=3, scale=2.0)
    R = synth.row_norm2_max(A)
    Ad = torch.as_tensor(A).cuda()
    a = SeqBasedSWFD(N, R, d, ell, schedule=schedule)
    a.fit(Ad[:n1])                                      # batched, n1 = 731: 11 rows wait in the open chunk
    sd = pickle.loads(pickle.dumps(a.state_dict()))     # what a checkpoint file would hold
    b = SeqBasedSWFD(N, R, d, ell, schedule=schedule)
    b.load_state_dict(sd)
    assert int(b.time) == int(a.time) == n1 and b.get_size() == a.get_size()
    for lo in range(n1, n2, 257):
        a.fit(Ad[lo:lo + 257])
        b.fit(Ad[lo:lo + 257])
    assert a.get_size() == b.get_size()
    _same_state(a, b)
    Ba, s2a, vta, da = a.get()
    Bb, s2b, vtb, db = b.get()
    assert a.last_level == b.last_level and da == db
    assert np.array_equal(Ba, Bb) and np.array_equal(s2a, s2b) and np.array_equal(vta, vtb)
    # the checkpoint itself is untouched by what followed: a third sketch restarts from it again
    c = SeqBasedSWFD(N, R, d, ell, schedule=schedule)
    c.load_state_dict(sd)
    c.fit(Ad[n1:n2])
    assert np.array_equal(c.get()[0], Ba)             # chunk boundaries follow the row count, not the calls


def test_multi_stream_state_and_configuration_check():
    from mused_b200 import synth
    from mused_b200.swfd import MultiStreamSWFD, SeqBasedSWFD
    K, d, ell, N = 3, 64, 16, 300
    X = np.stack([synth.lowrank_stream(900, d, seed=10 + k, scale=2.0) for k in range(K)])
    R = max(synth.row_norm2_max(X[k]) for k in range(K))
    Xd = torch.as_tensor(X).cuda()
    a = MultiStreamSWFD(K, N, R, d, ell)
    a.fit(Xd[:, :500])
    sd = a.state_dict()
    b = MultiStreamSWFD(K, N, R, d, ell)
    b.load_state_dict(sd)
    a.fit(Xd[:, 500:])
    b.fit(Xd[:, 500:])
    assert [a.get_size(k) for k in range(K)] == [b.get_size(k) for k in range(K)]
    for k in range(K):
        assert np.array_equal(a.get(k)[0], b.get(k)[0])
    other = SeqBasedSWFD(N, R, d, ell, schedule="batched")          # one stream: another allocation list
    with pytest.raises(ValueError):
        other.load_state_dict(sd)
    wrong = dict(sd, cfg=dict(other._cfg_key))                       # forged configuration: the library checks sizes itself
    with pytest.raises(ValueError):
        other.load_state_dict(wrong)
    ref = SeqBasedSWFD(N, R, d, ell)                                 # other schedule
    with pytest.raises(ValueError):
        ref.load_state_dict(dict(sd, cfg=dict(ref._cfg_key)))

"""GPU: the reference's Fast sketch family on the device operators (mused_b200/swfd/fast_swfd.py) against outputs of
the UNMODIFIED reference (tests/golden/fast_swfd.npz, oracle/gen_golden_fast.py): FastTimeBasedSWFD
(swfd/seq_based_swfd.py:209-311) -- B^T B of every query, get_size, queue lengths and (s, t) stamps of every
level -- and FastFrequentDirectionsWithDump (swfd/fastfdwithdump.py:7-138) -- the same rows dumped at the same
steps (compared through B^T B: singular vectors are sign-ambiguous), the same final stack and energy bound."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_fast_time_based_swfd_matches_reference(golden_dir):
    from mused_b200.swfd.fast_swfd import FastTimeBasedSWFD
    g = np.load(os.path.join(golden_dir, "fast_swfd.npz"))
    A, T, N, ell, qstep = g["A"], g["T"], int(g["N"]), int(g["ell"]), int(g["qstep"])
    sk = FastTimeBasedSWFD(N, float(g["R"]), A.shape[1], ell)
    assert sk.log_eNR == int(g["n_levels"])
    qi, pos = 0, 0
    for i in range(A.shape[0]):
        sk.fit(A[i:i + 1], int(T[i]))
        if (i + 1) % qstep == 0:
            B = np.asarray(sk.get()[0])
            ref = g["gram"][qi]
            assert np.abs(B.T @ B - ref).max() <= 1e-8 * np.abs(ref).max()
            assert sk.get_size() == int(g["size"][qi])
            ql = [len(sk.levels[j].fd.queue) for j in sk.levels] + [len(sk.levels[j].fd_aux.queue) for j in sk.levels]
            assert ql == list(g["qlen"][qi])
            st = [[int(e.s), int(e.t)] for j in sk.levels for e in sk.levels[j].fd.queue]
            n = int(g["stamps_split"][qi])
            assert st == g["stamps"][pos:pos + n].tolist()
            pos += n
            qi += 1
    assert qi == len(g["size"])


def test_fast_fd_with_dump_matches_reference(golden_dir):
    from mused_b200.swfd.fast_swfd import FastFrequentDirectionsWithDump
    g = np.load(os.path.join(golden_dir, "fast_swfd.npz"))
    A = g["A"]
    fd = FastFrequentDirectionsWithDump(A.shape[1], int(g["ell"]), float(g["fd_error"]))
    at, k = [], 0
    for i in range(300):
        r = fd.fit(A[i:i + 1])
        if r is not None:
            at.append(i)
            r = r.cpu().numpy()
            ref = g["fd_dump_gram"][k]
            assert np.abs(r.T @ r - ref).max() <= 1e-8 * np.abs(ref).max()
            k += 1
    assert at == list(g["fd_dump_at"])
    S = fd.get_sketch().cpu().numpy()
    assert S.shape[0] == int(g["fd_final_rows"])
    assert np.abs(S.T @ S - g["fd_final_gram"]).max() <= 1e-8 * np.abs(g["fd_final_gram"]).max()
    assert abs(fd.max_energy - float(g["fd_energy"])) <= 1e-9 * max(1.0, float(g["fd_energy"]))

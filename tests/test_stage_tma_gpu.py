"""K1, the TMA-staged append (k_bstage_tma: bulk loads / stores, one warp per row) against the SIMT append kernel it
replaces for contiguous fp32 rows: the engine state after the same stream must be BIT-identical (hi / lo planes and
the float64 ||x||^2 that decides routing are computed in the same order), for one and for several streams, ragged
row counts and calls below the TMA threshold.  The SIMT path is selected with MUSED_STAGE_SIMT=1 in a child process
(the switch is read once per process)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
from mused_b200 import synth
from mused_b200.swfd import SeqBasedSWFD, MultiStreamSWFD
out = {}
for d, ell, N, n in ((1024, 128, 3000, 7013), (96, 16, 400, 1501), (4096, 64, 1000, 2100)):
    A = synth.lowrank_stream(n, d, seed=d, scale=2.0)
    sk = SeqBasedSWFD(N, synth.row_norm2_max(A), d, ell, schedule="batched")
    Ad = torch.as_tensor(A).cuda()
    pos = 0
    for step in (1, 63, 64, 1000, 257, 5000, 10 ** 9):          # below and above the TMA threshold, ragged
        sk.fit(Ad[pos:pos + step]); pos += step
        if pos >= n: break
    B, s2, vt, delta = sk.get()
    out["B%%d" %% d], out["s%%d" %% d] = B, s2
    out["z%%d" %% d] = np.array([sk.get_size(), sk.last_level])
K, d, ell, N = 3, 256, 32, 500
X = np.stack([synth.lowrank_stream(1300, d, seed=k, scale=2.0) for k in range(K)])
ms = MultiStreamSWFD(K, N, max(synth.row_norm2_max(X[k]) for k in range(K)), d, ell)
ms.fit(torch.as_tensor(X).cuda())
for k in range(K):
    out["m%%d" %% k] = ms.get(k)[0]
np.savez(sys.argv[1], **out)
"""


def _run(tmp, simt):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ)
    env.pop("MUSED_STAGE_SIMT", None)
    if simt:
        env["MUSED_STAGE_SIMT"] = "1"
    path = os.path.join(tmp, "simt.npz" if simt else "tma.npz")
    p = subprocess.run([sys.executable, "-c", CHILD % root, path], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    return np.load(path)


def test_tma_staged_append_is_bit_identical_to_the_simt_append(tmp_path):
    a, b = _run(str(tmp_path), False), _run(str(tmp_path), True)
    assert sorted(a.files) == sorted(b.files) and len(a.files) == 12
    for k in a.files:
        assert np.array_equal(a[k], b[k]), k

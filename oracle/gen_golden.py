"""Generate tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference, imported through oracle/ref_shim.py) side by side with the
oracle restatement on seeded streams, asserting bit-identical outputs.

Run in the build container only:  python -m oracle.gen_golden
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import scipy

from . import ref_shim, swfd_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
VERSIONS = "numpy %s scipy %s" % (np.__version__, scipy.__version__)


def streams():
    rng = np.random.default_rng(1)
    A = synth.lowrank_stream(1500, 96, seed=0).astype(np.float32) * np.float32(1.5)
    yield "seq_lowrank", A, dict(N=300, ell=16, qstep=250)
    B = (synth.lowrank_stream(1200, 64, seed=1) * np.exp(rng.standard_normal(1200))[:, None] * 3).astype(np.float32)
    yield "seq_heavytail", B, dict(N=200, ell=8, qstep=200)
    W = 160
    Cm = (rng.random((3 * W, W)) < 0.06).astype(np.int64)
    yield "seq_adjacency01", Cm, dict(N=W, ell=10, qstep=W)
    D = synth.lowrank_stream(700, 160, seed=3).astype(np.float32) * np.float32(2.0)
    yield "seq_wide", D, dict(N=256, ell=48, qstep=175)


def run_seq(cls, A, N, ell, qstep, R):
    sk = cls(N, R, A.shape[1], ell)
    rec = []
    for t in range(A.shape[0]):
        sk.fit(A[t:t + 1].astype(np.float64) if A.dtype != np.int64 else A[t:t + 1])
        if (t + 1) % qstep == 0:
            size = sk.get_size()
            sketch, s2, vt, delta = sk.get()
            rec.append((t + 1, np.array(sketch), np.array(s2), np.array(vt), float(delta), size))
    return sk, rec


def main():
    swfd, mo, me = ref_shim.load()
    os.makedirs(OUT, exist_ok=True)
    for name, A, p in streams():
        A64 = A.astype(np.float64)
        R = float(np.max(np.einsum("ij,ij->i", A64, A64)))
        ref, rr = run_seq(swfd.SeqBasedSWFD, A, p["N"], p["ell"], p["qstep"], R)
        ora, ro = run_seq(so.SeqSWFD, A, p["N"], p["ell"], p["qstep"], R)
        for a, b in zip(rr, ro):
            assert a[0] == b[0] and a[5] == b[5] and a[4] == b[4]
            for x, y in zip(a[1:4], b[1:4]):
                assert np.array_equal(x, y), (name, a[0])
        st = ora.debug_state()
        # the oracle's final per-instance state must match the reference object tree
        for j in range(ref.logR):
            for side, sd in ((0, ref.levels[j].fd), (1, ref.levels[j].fd_aux)):
                o = st[2 * j + side]
                assert [int(e.s) for e in sd.queue] == o["q_s"] and [int(e.t) for e in sd.queue] == o["q_t"]
                assert int(sd.last_dump_time) == o["last_dump_time"]
                assert np.array_equal(np.asarray(sd.C.sigma_squared), o["s2"])
        out = dict(A=A, N=p["N"], ell=p["ell"], qstep=p["qstep"], R=R, versions=VERSIONS,
                   n_levels=ref.logR,
                   q_t=np.array([r[0] for r in rr]),
                   sigma2=np.stack([r[2] for r in rr]),
                   delta=np.array([r[4] for r in rr]),
                   size=np.array([r[5] for r in rr]),
                   gram=np.stack([r[1].T @ r[1] for r in rr]),
                   cov_err=np.array([so.cov_error(A64[max(0, r[0] - p["N"]):r[0]], r[1]) for r in rr]),
                   fin_last_dump=np.array([o["last_dump_time"] for o in st]),
                   fin_pending=np.array([o["pending"] for o in st]),
                   fin_energy=np.array([o["energy"] for o in st]),
                   fin_qlen=np.array([len(o["q_s"]) for o in st]),
                   fin_qs=np.concatenate([np.array(o["q_s"], dtype=np.int64) for o in st] + [np.zeros(0, np.int64)]),
                   fin_qt=np.concatenate([np.array(o["q_t"], dtype=np.int64) for o in st] + [np.zeros(0, np.int64)]))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, "levels", ref.logR, "queries", len(rr), "cov_err", out["cov_err"])

    # ---- single FD instance: FrequentDirectionsWithDump fit+dump loop ---------
    A = synth.lowrank_stream(400, 80, seed=5).astype(np.float32) * np.float32(2.0)
    A64 = A.astype(np.float64)
    err = 60.0
    rf = swfd.FrequentDirectionsWithDump(80, 12, err)
    of = so.FDDump(80, 12, err)
    dumps_t, dumps_v = [], []
    for t in range(A.shape[0]):
        rf.fit(A64[t:t + 1]); of.fit(A64[t:t + 1])
        v = rf.dump(); v2 = of.dump()
        assert (v is None) == (v2 is None)
        if v is not None:
            assert np.array_equal(v, v2)
            dumps_t.append(t); dumps_v.append(v[0])
    g1, g2 = rf.get(), of.get()
    for x, y in zip(g1[:3], g2[:3]):
        assert np.array_equal(x, y)
    np.savez_compressed(os.path.join(OUT, "fd_dump.npz"), A=A, ell=12, error=err, versions=VERSIONS,
                        dumps_t=np.array(dumps_t), dumps_v=np.stack([v * np.sign(v[np.argmax(np.abs(v))]) for v in dumps_v]),
                        sigma2=g1[1], gram=g1[0].T @ g1[0], delta=g1[3])
    print("fd_dump dumps", len(dumps_t))

    # ---- RobustFrequentDirections.fit(X, batch_size) --------------------------
    rr_ = swfd.RobustFrequentDirections(80, 12)
    oo_ = so.RFD(80, 12)
    rr_.fit(A64[:330], batch_size=12); oo_.fit(A64[:330], batch_size=12)
    rr_.fit(A64[330:], batch_size=5); oo_.fit(A64[330:], batch_size=5)
    g1, g2 = rr_.get(), oo_.get()
    for x, y in zip(g1[:3], g2[:3]):
        assert np.array_equal(x, y)
    assert g1[3] == g2[3]
    np.savez_compressed(os.path.join(OUT, "rfd_batch.npz"), A=A, ell=12, versions=VERSIONS,
                        sigma2=g1[1], gram=g1[0].T @ g1[0], delta=g1[3])
    print("rfd_batch delta", g1[3])


if __name__ == "__main__":
    main()

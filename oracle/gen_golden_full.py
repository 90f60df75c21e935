"""Full-length reference fixtures for BASELINE.json configs[0] and a long prefix of configs[1].

Runs the UNMODIFIED reference (/root/reference through oracle/ref_shim.py) row by row on the
seeded streams of the named shapes and stores, per query, what a parity test needs without the
full sketches: sigma^2, delta, get_size(), chosen level, cov-error of the window
(swfd/test_synthetic.py:83-92), ||B^T B||_F, trace and the projected Gram P^T B^T B P on a seeded
d x 32 probe (sign-free), plus the final snapshot (s, t) lists of every FD instance.  configs[0]
also stores the cosine-threshold event assignment of every window (oracle/ops_oracle.cosine_assign)
and the reference's own metrics_evaluation.compute_all_metrics scores of those labels.  The oracle
restatement is run on the same rows and asserted bit-identical to the reference (--no-oracle skips).

Build container only (needs /root/reference):
    OMP_NUM_THREADS=1 python -m oracle.gen_golden_full c0
    OMP_NUM_THREADS=1 python -m oracle.gen_golden_full c1 [--rows 12000]
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import argparse
import hashlib
import os
import sys
import time

import numpy as np
import scipy
from scipy import linalg

from . import ref_shim, swfd_oracle as so, ops_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mused_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
VERSIONS = "numpy %s scipy %s" % (np.__version__, scipy.__version__)
PROBE_COLS = 32


def probe(d):
    return np.random.default_rng([d, 4242]).standard_normal((d, PROBE_COLS)) / np.sqrt(d)


def sha(A):
    return hashlib.sha256(np.ascontiguousarray(A).tobytes()).hexdigest()


def config(name, rows=None):
    if name == "c0":      # configs[0]: 20k x 512 multimodal event stream, l=64, N=2000
        n = rows or 20000
        A, labels, centres = synth.event_stream(n, blocks=(256, 192, 32, 32), n_events=150, seed=0)
        return dict(A=A, labels=labels, centres=centres, N=2000, ell=64, qstep=2000, R=synth.row_norm2_max(A), tau=0.5)
    if name == "c1":      # configs[1] prefix: d=1024, l=128, N=10000 low-rank + noise, R bound 64 (bench.py)
        n = rows or 12000
        A = synth.lowrank_stream(n, 1024, seed=0)
        assert synth.row_norm2_max(A) <= 64.0
        return dict(A=A, labels=None, centres=None, N=10000, ell=128, qstep=2500, R=64.0, tau=None)
    if name in ("l256", "l512"):   # strict engine beyond sketch_dim 128 (configs[3] / configs[4] sketch sizes, small d)
        n, d, seed, N, ell = (2000, 640, 7, 1500, 256) if name == "l256" else (1500, 768, 8, 1200, 512)
        A = synth.lowrank_stream(rows or n, d, seed=seed, scale=2.0)
        return dict(A=A, labels=None, centres=None, N=N, ell=ell, qstep=500, R=synth.row_norm2_max(A), tau=None)
    raise SystemExit("unknown config " + name)


def ref_level(ref):
    """seq_based_swfd.py:83-98 level choice, recomputed without touching state."""
    rj = 0
    for j in range(ref.logR):
        q = ref.levels[j].fd.queue
        if len(q) == 0:
            break
        if int(ref.time) - int(q[0].s) >= min(ref.N - 1, int(ref.time) - 1):
            rj = j
            break
    return rj


def summarise(B, s2, delta, size, level, A_w, P):
    BP = B @ P
    G = B @ B.T
    return dict(sigma2=np.array(s2), delta=float(delta), size=int(size), level=int(level),
                cov_err=so.cov_error(A_w, B), fro=float(np.linalg.norm(G)), trace=float(np.trace(G)),
                pgram=BP.T @ BP)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["c0", "c1", "l256", "l512"])
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--no-oracle", action="store_true")
    args = ap.parse_args()
    cfg = config(args.config, args.rows or None)
    A, N, ell, qstep, R = cfg["A"], cfg["N"], cfg["ell"], cfg["qstep"], cfg["R"]
    n, d = A.shape
    A64 = A.astype(np.float64)
    P = probe(d)
    swfd, mo, me = ref_shim.load()
    ref = swfd.SeqBasedSWFD(N, R, d, ell)
    ora = None if args.no_oracle else so.SeqSWFD(N, R, d, ell)
    recs, labels_all = [], []
    t0 = time.perf_counter()
    t_ref = 0.0
    for t in range(n):
        x = A64[t:t + 1]
        ta = time.perf_counter()
        ref.fit(x)
        t_ref += time.perf_counter() - ta
        if ora is not None:
            ora.fit(x)
        if (t + 1) % qstep == 0 or t + 1 == n:
            lvl = ref_level(ref)
            size = ref.get_size()
            B, s2, vt, delta = ref.get()
            if ora is not None:
                assert ora.pick_level() == lvl and ora.get_size() == size
                Bo, s2o, vto, do = ora.get()
                assert np.array_equal(B, Bo) and np.array_equal(s2, s2o) and do == delta, (args.config, t + 1)
            A_w = A64[max(0, t + 1 - N):t + 1]
            r = summarise(np.array(B), s2, delta, size, lvl, A_w, P)
            r["t"] = t + 1
            recs.append(r)
            print("%s t=%d level=%d size=%d cov_err=%.6f (bound 1/l=%.6f) delta=%.4f  %.0fs (ref fit %.1f rows/s)"
                  % (args.config, t + 1, lvl, size, r["cov_err"], 1.0 / ell, delta, time.perf_counter() - t0,
                     (t + 1) / t_ref), flush=True)
    out = dict(config=args.config, n=n, d=d, N=N, ell=ell, qstep=qstep, R=R, versions=VERSIONS, sha_A=sha(A),
               n_levels=ref.logR, ref_fit_rows_per_s=n / t_ref,
               q_t=np.array([r["t"] for r in recs]),
               sigma2=np.stack([r["sigma2"] for r in recs]),
               delta=np.array([r["delta"] for r in recs]),
               size=np.array([r["size"] for r in recs]),
               level=np.array([r["level"] for r in recs]),
               cov_err=np.array([r["cov_err"] for r in recs]),
               fro=np.array([r["fro"] for r in recs]),
               trace=np.array([r["trace"] for r in recs]),
               pgram=np.stack([r["pgram"] for r in recs]))
    # final snapshot queues of every FD instance, straight from the reference's object tree
    qlen, qs, qt, ld, s2f = [], [], [], [], []
    for j in range(ref.logR):
        for sd in (ref.levels[j].fd, ref.levels[j].fd_aux):
            qlen.append(len(sd.queue))
            qs += [int(e.s) for e in sd.queue]
            qt += [int(e.t) for e in sd.queue]
            ld.append(int(sd.last_dump_time))
            s2f.append(np.asarray(sd.C.sigma_squared, dtype=np.float64))
    out.update(fin_qlen=np.array(qlen), fin_qs=np.array(qs, dtype=np.int64), fin_qt=np.array(qt, dtype=np.int64),
               fin_last_dump=np.array(ld, dtype=np.int64), fin_s2=np.stack(s2f))
    if cfg["labels"] is not None:
        # configs[0]: cosine-threshold event assignment of every window against the event centres and the
        # reference's own scorer (metrics_evaluation.py:36-102) on the whole stream
        C = cfg["centres"].astype(np.float64)
        lab, score, gap = ops_oracle.cosine_assign(A64, C, cfg["tau"])
        clusters = (lab + 1).astype(np.int64)              # -1 (below tau) -> 0 = noise label of main.py:306
        res, _ = me.get_initial_results()
        me.compute_all_metrics(res, n, 0.95, "multi", "none", ell, 50, N, list(clusters), list(cfg["labels"]), 1, 0)
        out.update(tau=cfg["tau"], assign_labels=lab.astype(np.int32), assign_margin=gap.astype(np.float32),
                   assign_score=score.astype(np.float32), true_labels=cfg["labels"].astype(np.int32),
                   nmi=res["nmi_score"][0], nmi_e=res["nmi_e_score"][0], f1=res["f1_score"][0],
                   precision=res["precision"][0], recall=res["recall"][0], accuracy=res["accuracy"][0])
        print("assign: nmi %.6f nmi_e %.6f f1 %.6f" % (out["nmi"], out["nmi_e"], out["f1"]))
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, "full_%s.npz" % args.config)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; reference fit %.1f rows/s" % (n / t_ref))


if __name__ == "__main__":
    main()

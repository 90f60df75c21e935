#!/usr/bin/env python
"""bench.py -- stream rows/sec of the SWFD update + assign hot path.

Workload (BASELINE.json configs[1]): d=1024, l=128, window N=10k synthetic
stream (seeded low-rank + noise model of swfd/generate_synthetic.py).  A STEP is
one window of 10 000 rows: SeqBasedSWFD.fit of the rows (reference flush / dump
schedule, float64), one SeqBasedSWFD.get() at the window boundary, and the fused
cosine+threshold+argmax assignment of the window's rows against 150 centroids.

  python bench.py [--gpus N] [--steps K] [--warmup W]          # B200 arm
  python bench.py --impl reference ...                          # CPU arm (oracle port)

N > 1 (torchrun): `value` = every rank runs an independent stream (the path shards by
independent streams/windows; no data-path collective), rows of all ranks / max-over-ranks
time, scaling "weak" -- the same metric and config as N = 1.  The line also carries
`sharded`: north_star's column-sharded mode (d split over the N ranks, one NCCL allreduce of
the Gram per shrink, issued from inside the C ABI on the engine's stream) on d = 16384,
l = 128, the same stream on every rank, "scaling": "strong"; at N = 1 the same key holds the
unsharded single-GPU figure of that shape.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D, ELL, WIN = 1024, 128, 10000
K_CENTROIDS, TAU = 150, 0.2
R_BOUND = 64.0      # upper bound of ||row||^2 handed to SeqBasedSWFD (6 levels); checked on the data
METRIC = "stream rows/sec (SWFD update+assign) at d=1024,l=128; cov-error vs reference"
UNIT = "rows/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cluster", type=int, default=0)
    ap.add_argument("--cpu-rows", type=int, default=0, help="rows of the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--streams", default="4,16", help="comma list: independent streams batched per GPU by the native "
                    "many-streams engine in the multi_stream leg ('' = skip)")
    ap.add_argument("--no-sharded", action="store_true", help="skip the column-sharded d=16384 leg")
    ap.add_argument("--phase-times", action="store_true", help="diagnostic: synchronise after assign and print fit/get/assign wall ms")
    ap.add_argument("--schedule", default="batched", choices=["batched", "reference"],
                    help="batched: north_star's chunked tensor-core schedule (headline); "
                         "reference: the reference's per-row flush schedule")
    ap.add_argument("--no-refsched", action="store_true", help="skip the extra reference-schedule timing")
    ap.add_argument("--config", type=int, default=1, choices=[1, 3],
                    help="BASELINE.json configs[] index of the workload: 1 = the metric's own (d=1024, l=128, N=10k); "
                         "3 = d=4096, l=256, N=100k (tensor-bound Gram / rotate; diagnostic line, device legs only)")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return float(j["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def measured_tensor_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["bf16_tflops"])
    return 1590.0


def make_stream(n_rows, seed):
    from mused_b200 import synth
    A = synth.lowrank_stream(n_rows, D, seed=seed)
    assert synth.row_norm2_max(A) <= R_BOUND
    return A, R_BOUND


def centroids(seed):
    rng = np.random.default_rng([seed, 99])
    C = rng.standard_normal((K_CENTROIDS, D)).astype(np.float32)
    return C / np.linalg.norm(C, axis=1, keepdims=True)


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (the profiling recipe's
    clocks line), read through NVML in-process every 250 ms: a polling nvidia-smi
    child process slowed this launch-heavy step by ~40 % (it serialises with kernel
    launches in the driver), NVML queries from a thread do not."""

    def __init__(self, index):
        self.index = index
        self.sm, self.mx, self.reasons = [], [], set()
        self.stop_flag = threading.Event()
        self.thread = None
        self.err = None

    def _run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if vis:
                try:
                    idx = int(vis.split(",")[self.index])
                except Exception:
                    idx = self.index
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                     "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                     "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                     "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
            while not self.stop_flag.is_set():
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                self.mx.append(mx)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for k, bit in names.items():
                        if r & bit:
                            self.reasons.add(k)
                except Exception:
                    pass
                self.stop_flag.wait(0.25)
        except Exception as e:                      # pragma: no cover
            self.err = repr(e)

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        if self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["not sampled"]}
        self.stop_flag.set()
        self.thread.join(timeout=2.0)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: %s" % self.err]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": max(self.mx), "reasons": sorted(self.reasons),
                "samples": len(self.sm), "source": "nvml"}


# ---------------------------------------------------------------------------
def cpu_port_rate(A, R, C, warm_rows, rows, with_query=True):
    """rows/s of the oracle port (NumPy/SciPy restatement of the reference) on
    the host cores: rows [warm_rows, warm_rows+rows) after an untimed warm-up."""
    from oracle import swfd_oracle as so, ops_oracle
    sk = so.SeqSWFD(WIN, R, D, ELL)
    A64 = A[:warm_rows + rows].astype(np.float64)
    for t in range(warm_rows):
        sk.fit(A64[t:t + 1])
    t0 = time.perf_counter()
    for t in range(warm_rows, warm_rows + rows):
        sk.fit(A64[t:t + 1])
    if with_query:
        sk.get()
        ops_oracle.cosine_assign(A64[warm_rows:warm_rows + rows], C, TAU)
    dt = time.perf_counter() - t0
    return rows / dt, dt


def blas_limit(n):
    from threadpoolctl import threadpool_limits
    return threadpool_limits(limits=n, user_api="blas")


def calibrate_blas_threads(A, R, rows=16):
    """The port's hot call is a LAPACK gesvd of a (l + b) x d buffer, b small: OpenBLAS threads can hurt (round 1:
    16 threads 13.3 rows/s, 1 thread 21.6 rows/s).  Time `rows` rows with 1 thread and with all, keep the faster."""
    from oracle import swfd_oracle as so
    best = (None, 0.0)
    A64 = A[:64 + 2 * rows].astype(np.float64)
    for n in (1, os.cpu_count() or 1):
        with blas_limit(n):
            sk = so.SeqSWFD(WIN, R, D, ELL)
            for t in range(64):
                sk.fit(A64[t:t + 1])
            t0 = time.perf_counter()
            for t in range(64, 64 + rows):
                sk.fit(A64[t:t + 1])
            rate = rows / (time.perf_counter() - t0)
        if rate > best[1]:
            best = (n, rate)
    return best[0]


def cpu_batched_rate(A, R, rows):
    """rows/s of oracle/swfd_batched_oracle.py (the float64 statement of the batched schedule) on the host."""
    from oracle.swfd_batched_oracle import BatchedSeqSWFD
    sk = BatchedSeqSWFD(WIN, R, D, ELL)
    A64 = A[:ELL + rows].astype(np.float64)
    sk.fit(A64[:ELL])
    t0 = time.perf_counter()
    sk.fit(A64[ELL:ELL + rows])
    dt = time.perf_counter() - t0
    return rows / dt, dt


def host_threads():
    try:
        from threadpoolctl import threadpool_info
        n = [p.get("num_threads", 1) for p in threadpool_info() if p.get("user_api") == "blas"]
        return max(n) if n else 1
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """CPU arm: the reference's algorithm (oracle port; the Python reference
    itself cannot travel to the GPU box) on the host cores, bounded samples."""
    if rank != 0:
        return
    rows_per_step = 48
    warm_fill = 160                      # untimed: sketches leave the all-zero start
    total = warm_fill + (args.warmup + args.steps) * rows_per_step
    A, R = make_stream(max(total, 2048), seed=0)
    C = centroids(0)
    from oracle import swfd_oracle as so, ops_oracle
    nthr = calibrate_blas_threads(A, R)
    ctx = blas_limit(nthr)
    ctx.__enter__()
    sk = so.SeqSWFD(WIN, R, D, ELL)
    A64 = A[:total].astype(np.float64)
    for t in range(warm_fill):
        sk.fit(A64[t:t + 1])
    pos = warm_fill
    times = []
    for s in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        for t in range(pos, pos + rows_per_step):
            sk.fit(A64[t:t + 1])
        ops_oracle.cosine_assign(A64[pos:pos + rows_per_step], C, TAU)
        times.append(time.perf_counter() - t0)
        pos += rows_per_step
    tq = time.perf_counter()
    sk.get()
    tq = time.perf_counter() - tq
    timed = times[args.warmup:]
    # one get() per window of WIN rows, amortised like the B200 arm does
    sec_per_row = sum(timed) / (len(timed) * rows_per_step) + tq / WIN
    value = 1.0 / sec_per_row
    ctx.__exit__(None, None, None)
    cores = nthr
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(timed) / len(timed),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "SWFD d=1024 l=128 N=10000 low-rank+noise stream + cosine assign k=150 (configs[1])",
                   "schedule": "reference", "levels": sk.logR},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d timed steps of %d rows after %d warm-up rows; one get() of %.3fs amortised over %d rows; "
                                   "BLAS threads = %d (the faster of 1 and %d on a 16-row calibration)"
                                   % (args.steps, rows_per_step, warm_fill + args.warmup * rows_per_step, tq, WIN,
                                      nthr, os.cpu_count() or 1)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from mused_b200 import _lib
    from mused_b200.swfd import SeqBasedSWFD
    from mused_b200 import matrix_operations as mo

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    n_steps = args.warmup + args.steps
    A, R = make_stream(n_steps * WIN, seed=rank)          # independent stream per rank
    Cn = centroids(rank)
    Cd = torch.as_tensor(Cn).to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    Ad = torch.as_tensor(A).to(dev)                        # inputs resident in HBM before timing
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def device_arm(schedule, warmup, steps, sample_clocks):
        """`value`: whole windows (fit + get + assign) with inputs resident in HBM."""
        sk = SeqBasedSWFD(WIN, R, D, ELL, device=local_rank, cluster=args.cluster, schedule=schedule)

        get_s = [0.0, 0.0, 0.0]                             # host wall seconds: get, fit (to sync), assign

        def step(i):
            blk = Ad[i * WIN:(i + 1) * WIN]
            tf = time.perf_counter()
            sk.fit(blk)
            torch.cuda.current_stream().synchronize()      # get() starts with this wait; time it apart
            tg = time.perf_counter()
            out = sk.get()
            ta = time.perf_counter()
            lab = mo.cosine_assign(blk, Cd, TAU, device=dev)
            if args.phase_times:
                torch.cuda.current_stream().synchronize()
            get_s[0] += ta - tg
            get_s[1] += tg - tf
            get_s[2] += time.perf_counter() - ta
            return out, lab

        for i in range(warmup):
            step(i)
        barrier()
        sampler = ClockSampler(local_rank)
        if sample_clocks and rank == 0:
            sampler.start()
        if schedule == "batched":
            sk.kernel_ms(enable=1)
        else:
            sk.kernel_timing(enable=1)
        l0 = lib.mused_launch_count()
        get_s[0] = get_s[1] = get_s[2] = 0.0
        e0.record()
        for i in range(warmup, warmup + steps):
            out, lab = step(i)
        e1.record()
        barrier()
        ms_total = max_over_ranks(e0.elapsed_time(e1))
        res = {"ms_total": ms_total, "launches": int(lib.mused_launch_count() - l0),
               "value": world * steps * WIN / (ms_total * 1e-3), "sketch": out[0],
               "get_ms": 1e3 * get_s[0] / steps, "fit_ms": 1e3 * get_s[1] / steps, "assign_ms": 1e3 * get_s[2] / steps,
               "clocks": sampler.stop() if (sample_clocks and rank == 0) else None}
        if schedule == "batched":
            res["kernel_ms"], res["chunks"], res["k_rows"] = sk.kernel_ms(enable=0)
            res["gaps_ms"] = sk.kernel_gaps()
        else:
            res["k_ms"], res["k_launches"], res["k_rows"] = sk.kernel_timing(enable=0)
            res["prof"] = sk.profile()
        res["state"] = sk.debug_state()
        res["level"] = sk.last_level
        sk.close()
        return res

    main_res = device_arm(args.schedule, args.warmup, args.steps, True)
    other = None
    if args.schedule == "batched" and not args.no_refsched:
        # the bit-faithful reference flush schedule on the same stream (1 warm-up + 2 timed windows), device-resident
        # rows, then one more window from HOST buffers through the same public calls (its e2e)
        other = device_arm("reference", 1, min(2, args.steps), False)
        sk3 = SeqBasedSWFD(WIN, R, D, ELL, device=local_rank, cluster=args.cluster, schedule="reference")
        sk3.fit(A[:WIN])
        sk3.get()
        lab_h = torch.empty(WIN, dtype=torch.int32).pin_memory()
        barrier()
        e0.record()
        sk3.fit(A[WIN:2 * WIN])                        # numpy host rows: H2D inside the C-ABI call
        sk3.get()
        lab_h.copy_(mo.cosine_assign(torch.as_tensor(A[WIN:2 * WIN]).to(dev), Cd, TAU, device=dev))
        e1.record()
        barrier()
        other["e2e_value"] = world * WIN / (max_over_ranks(e0.elapsed_time(e1)) * 1e-3)
        sk3.close()

    # ---------------- independent streams batched on one GPU (north_star: "independent streams and windows are
    # batched per GPU"; configs[4]: "64 independent windows batched per GPU"): the native many-streams engine
    # (mused_swfd_cfg.n_streams): ONE K1..K5 launch sequence per chunk covers K x 2 x levels FD instances
    multi = None
    ks = [int(x) for x in str(args.streams).split(",") if x.strip()] if args.schedule == "batched" else []
    if ks and n_steps >= 3:
        from mused_b200.swfd import MultiStreamSWFD
        multi = {"unit": UNIT, "api": "MultiStreamSWFD / mused_swfd_fit_streams (n_streams handles, one launch sequence)",
                 "step": "fit of one window per stream + get() of every stream + assign of every stream's window",
                 "by_streams": {}}
        for K in ks:
            wins = 2
            # K x rows x d, resident before timing; streams cycle through this rank's windows at different offsets
            Ak = torch.stack([Ad[((k % n_steps) * WIN):((k % n_steps) * WIN) + WIN] for k in range(K)])
            Aw = [torch.stack([Ad[(((k + w) % n_steps) * WIN):(((k + w) % n_steps) * WIN) + WIN] for k in range(K)])
                  for w in range(1, 1 + wins)]
            msk = MultiStreamSWFD(K, WIN, R, D, ELL, device=local_rank)
            msk.fit(Ak)
            for k in range(K):
                msk.get(k)
            barrier()
            e0.record()
            for w in range(wins):
                msk.fit(Aw[w])
                for k in range(K):
                    msk.get(k)
                mo.cosine_assign(Aw[w].reshape(K * WIN, D), Cd, TAU, device=dev)
            e1.record()
            barrier()
            ms = max_over_ranks(e0.elapsed_time(e1))
            multi["by_streams"][str(K)] = {"value": world * K * wins * WIN / (ms * 1e-3), "ms_per_window": ms / wins}
            msk.close()
            del Ak, Aw, msk
            torch.cuda.empty_cache()
        best = max(multi["by_streams"].items(), key=lambda kv: kv[1]["value"])
        multi["streams_per_gpu"], multi["value"] = int(best[0]), best[1]["value"]

    # ---------------- column-sharded mode (north_star): d = 16384 split over the `world` ranks, the same
    # rows on every rank, per chunk two NCCL allreduces (||x||^2, Gram) posted from inside the C ABI on the
    # engine's stream; strong scaling: the same total work for every N (N = 1: the unsharded engine)
    sharded = None
    if args.schedule == "batched" and not args.no_sharded:
        from mused_b200.sharded import NcclShardedSWFD, shard_bounds
        DS, wins = 16384, 2
        gen = torch.Generator(device=dev)
        gen.manual_seed(1234)                                  # the same stream on every rank
        Us, _ = torch.linalg.qr(torch.randn(DS, 10, device=dev, generator=gen))
        Dg = torch.tensor([1.0 - (i - 1) / 10 for i in range(10)], device=dev)
        lo, hi = shard_bounds(DS, world, rank)
        Xs = []
        for w in range(1 + wins):                              # full rows are drawn, this rank keeps its columns
            S_ = torch.randn(WIN, 10, device=dev, generator=gen) * Dg
            Z_ = torch.randn(WIN, DS, device=dev, generator=gen)
            Xw = S_ @ Us.T + Z_ * (1.0 / 40.0)                 # zeta scaled by sqrt(d / 1024): same ||row||^2 as configs[1]
            Xs.append(Xw[:, lo:hi].contiguous())
            del S_, Z_, Xw
        shd = NcclShardedSWFD(WIN, R, DS, ELL, device=local_rank)
        shd.fit(Xs[0])
        shd.get(gather=False)
        barrier()
        e0.record()
        for w in range(1, 1 + wins):
            shd.fit(Xs[w])
            shd.get(gather=False)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        na, nb = shd.comm_stats()
        sharded = {"value": wins * WIN / (ms * 1e-3), "unit": UNIT, "scaling": "strong", "n_gpus": world,
                   "workload": "SWFD d=16384 l=128 N=10000 low-rank+noise stream, d column-sharded over %d GPU(s), "
                               "fit + get per window (sketch left sharded)" % world,
                   "columns_per_gpu": hi - lo, "ms_per_window": ms / wins,
                   "allreduces_per_window": na / (1 + wins), "allreduce_MB_per_window": nb / (1 + wins) / 1e6,
                   "collective": "ncclAllReduce(float64 sum) of ||x||^2[rows] and of the Gram [S][2l][2l] per chunk, "
                                 "posted from the C ABI on the engine's stream"}
        shd.close()
        del Xs, shd
        torch.cuda.empty_cache()

    # ---------------- end-to-end arm (`e2e`): host buffers through the public API
    e2e = None
    if not args.no_e2e:
        Ah = torch.as_tensor(A).pin_memory()
        An = Ah.numpy()
        sk2 = SeqBasedSWFD(WIN, R, D, ELL, device=local_rank, cluster=args.cluster, schedule=args.schedule)
        lab_host = torch.empty(WIN, dtype=torch.int32).pin_memory()

        get2_s = [0.0]

        def step_e2e(i):
            blk = An[i * WIN:(i + 1) * WIN]
            sk2.fit(blk)                                   # H2D inside the C-ABI call
            torch.cuda.current_stream().synchronize()
            tg = time.perf_counter()
            out = sk2.get()                                # sketch / sigma^2 / Vt D2H inside
            get2_s[0] += time.perf_counter() - tg
            xd = Ah[i * WIN:(i + 1) * WIN].to(dev, non_blocking=True)
            lab = mo.cosine_assign(xd, Cd, TAU, device=dev)
            lab_host.copy_(lab, non_blocking=True)
            torch.cuda.synchronize()
            return out

        for i in range(args.warmup):
            step_e2e(i)
        barrier()
        t0 = time.perf_counter()
        get2_s[0] = 0.0
        e0.record()
        for i in range(args.warmup, n_steps):
            step_e2e(i)
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
        e2e = {"value": world * args.steps * WIN / (ms_e2e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": 2 * WIN * D * 4,      # rows for fit + rows for assign
               "d2h_bytes_per_step": 2 * ELL * D * 8 + ELL * 8 + 8 + WIN * 4,
               "wall_s": time.perf_counter() - t0, "get_ms_per_step": 1e3 * get2_s[0] / args.steps}
        sk2.close()
    del Ad

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel + per-kernel figures ----------
    S = len(main_res["state"])
    bytes_per_row = 8 * D + S * 20 * D                     # append 8d + S*20d per row
    peak, how = measured_peaks()
    tpeak = measured_tensor_peak()
    roofline, kernels = None, None
    if args.schedule == "batched":
        km, chunks = main_res["kernel_ms"], max(1, main_res["chunks"])
        rows_per_launch = main_res["k_rows"] / chunks
        dom = max(km, key=km.get)
        dur_s = km[dom] / chunks * 1e-3
        achieved = bytes_per_row * rows_per_launch / dur_s / 1e9
        # dram bytes / issue figures of the dominant kernel come from the committed ncu capture of THIS command
        # (profiles/r02_ncu_dominant.json, written by tools/ncu_summarise.py); absent -> null, never a pasted constant
        ncu = {}
        pj = os.path.join(ROOT, "profiles", "r02_ncu_dominant.json")
        if os.path.exists(pj):
            with open(pj) as f:
                ncu = json.load(f)
        kname = {"K1_stage": "k_bstage", "K5_book": "k_bbook", "K2_gram": "k_gram_tc" if ELL <= 128 else "k_gram_tiled",
                 "K3_eig": "k_eig_jacobi", "K4_rotate": "k_rotate_tc"}[dom]
        if args.config != 1:
            ncu = {}                                       # the committed ncu figures are of the configs[1] capture
        nk = ncu.get(kname, {})
        roofline = {"bound": "hbm", "kernel": kname,
                    "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": nk.get("dram_bytes_per_launch"),
                    "ncu": {k: v for k, v in nk.items() if k != "dram_bytes_per_launch"} or None,
                    "peak_source": how, "algorithmic_bytes_per_row": bytes_per_row,
                    "rows_per_launch": rows_per_launch, "avg_launch_ms": km[dom] / chunks,
                    "share_of_step": km[dom] / main_res["ms_total"],
                    "note": "dominant kernel is the batched mixed-precision Jacobi eigensolve (shared-memory resident, 8 warps "
                            "x 2-column register blocks per CTA, cluster of 8 CTAs per FD instance): a sweep is 255 dependent "
                            "rotation steps of ~600 cycles each (dot -> 5 shuffle stages -> angle with 2 MUFU -> broadcast -> "
                            "rotate), ~10 sweeps per chunk -- a dependent-issue chain, no HBM or tensor roofline applies; "
                            "bytes are SURVEY 8(d)'s per-row figure (DESIGN.md section 4, cycle accounting in profiles/)"}
        n2 = 2 * ELL
        us = {k: v / chunks * 1e3 for k, v in km.items()}
        kernels = {
            "K1_stage": {"us_per_chunk": us["K1_stage"], "bound": "hbm",
                         "achieved_GBs": rows_per_launch * D * 12 / (us["K1_stage"] * 1e-6) / 1e9,
                         "algorithmic_bytes": "rows*d*(4 read + 8 hi/lo write)"},
            "K2_gram": {"us_per_chunk": us["K2_gram"], "bound": "tensor",
                        "achieved_TFLOPs_dense": S * 2.0 * n2 * n2 * D / (us["K2_gram"] * 1e-6) / 1e12,
                        "issued_TFLOPs_3xTF32": 3 * S * 2.0 * n2 * n2 * D / (us["K2_gram"] * 1e-6) / 1e12,
                        "peak_bf16_TFLOPs": tpeak, "note": "kind::tf32 runs at half the bf16 rate; 3 passes"},
            "K3_eig": {"us_per_chunk": us["K3_eig"], "bound": "dependent-issue chain / cluster barrier",
                       "share_of_step": km["K3_eig"] / main_res["ms_total"]},
            "K4_rotate": {"us_per_chunk": us["K4_rotate"], "bound": "tensor",
                          "achieved_TFLOPs_dense": S * 2.0 * ELL * n2 * D / (us["K4_rotate"] * 1e-6) / 1e12,
                          "issued_TFLOPs_3xTF32": 3 * S * 2.0 * ELL * n2 * D / (us["K4_rotate"] * 1e-6) / 1e12,
                          "peak_bf16_TFLOPs": tpeak},
            "K5_book": {"us_per_chunk": us["K5_book"]},
            "stream_idle": {"us_per_chunk_inside": main_res["gaps_ms"][0] / chunks * 1e3,
                            "us_per_chunk_between": main_res["gaps_ms"][1] / chunks * 1e3,
                            "note": "CUDA-event time between the end of one kernel and the start of the next"},
        }
    elif main_res.get("k_launches"):
        rows_per_launch = main_res["k_rows"] / main_res["k_launches"]
        dur_s = (main_res["k_ms"] / main_res["k_launches"]) * 1e-3
        achieved = bytes_per_row * rows_per_launch / dur_s / 1e9
        roofline = {"bound": "hbm", "kernel": "k_swfd_rows", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": None, "peak_source": how,
                    "algorithmic_bytes_per_row": bytes_per_row, "rows_per_launch": rows_per_launch,
                    "avg_launch_ms": main_res["k_ms"] / main_res["k_launches"],
                    "share_of_step": main_res["k_ms"] / main_res["ms_total"],
                    "note": "reference flush schedule is a per-row dependency chain (latency/FP64-bound), see DESIGN.md"}

    # covariance error ON ONE WINDOW for all three: rows 0..10000 of the fixture stream of tests/golden/full_c1.npz
    # (generated offline by the UNMODIFIED reference, oracle/gen_golden_full.py), queries at the fixture's times
    from oracle import swfd_oracle as so
    cov = {"bound_1_over_l": 1.0 / ELL,
           "last_timed_window_" + args.schedule: so.cov_error(A[(n_steps - 1) * WIN:n_steps * WIN].astype(np.float64),
                                                               main_res["sketch"])}
    fx = os.path.join(ROOT, "tests", "golden", "full_c1.npz")
    if os.path.exists(fx) and args.config == 1:
        import hashlib
        from mused_b200 import synth
        g = np.load(fx)
        Af = synth.lowrank_stream(int(g["n"]), D, seed=0)
        if hashlib.sha256(np.ascontiguousarray(Af).tobytes()).hexdigest() == str(g["sha_A"]):
            qts = [int(t) for t in g["q_t"] if int(t) <= WIN]
            Afd = torch.as_tensor(Af).to(dev)
            errs = {}
            for sched in ("reference", "batched"):
                skf = SeqBasedSWFD(WIN, R, D, ELL, device=local_rank, schedule=sched)
                t0_ = 0
                for t in qts:
                    skf.fit(Afd[t0_:t])
                    t0_ = t
                    Bf = skf.get()[0]
                errs[sched] = so.cov_error(Af[max(0, qts[-1] - WIN):qts[-1]].astype(np.float64), Bf)
                skf.close()
            qi = list(g["q_t"]).index(qts[-1])
            cov.update({"window": "rows %d..%d of the fixture stream (tests/golden/full_c1.npz), queries at %s"
                                  % (max(0, qts[-1] - WIN), qts[-1], qts),
                        "reference_cpu": float(g["cov_err"][qi]),
                        "reference_cpu_source": "imported reference, generated offline (%s)" % str(g["versions"]),
                        "reference_schedule_gpu": errs["reference"], "batched_gpu": errs["batched"],
                        "rel_diff_reference_schedule": abs(errs["reference"] - float(g["cov_err"][qi])) / float(g["cov_err"][qi])})
        else:
            cov["window"] = "fixture stream mismatch: same-window comparison skipped"

    cpu, cpu_b = None, None
    if not args.no_cpu:
        rows = args.cpu_rows or 384
        nthr = calibrate_blas_threads(A, R)
        with blas_limit(nthr):
            v, dt = cpu_port_rate(A, R, Cn, warm_rows=384, rows=rows)
            vb, dtb = cpu_batched_rate(A, R, rows=8 * ELL)
        cpu = {"value": v, "unit": UNIT, "cores": nthr, "kind": "port",
               "sample": "rows 384..%d of the same stream (after 384 untimed warm-up rows) + one get() + assign, %.1fs; "
                         "BLAS threads = %d (the faster of 1 and %d on a 16-row calibration)"
                         % (384 + rows, dt, nthr, os.cpu_count() or 1)}
        cpu_b = {"value": vb, "unit": UNIT, "cores": nthr, "kind": "port (oracle/swfd_batched_oracle.py, the float64 "
                 "statement of the batched schedule)", "sample": "%d rows after %d warm-up rows, %.1fs" % (8 * ELL, ELL, dtb)}

    st = main_res["state"]
    line = {
        "metric": METRIC, "value": main_res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": main_res["ms_total"] / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 rows, 3xTF32 products, f32+f64 eigensolve"
        if args.schedule == "batched" else "f64", "data": "synthetic",
        "config": {"workload": "SWFD d=%d l=%d N=%d low-rank+noise stream + cosine assign k=150 (configs[%d])"
                               % (D, ELL, WIN, args.config),
                   "schedule": args.schedule, "levels": S // 2, "fd_instances": S, "rows_per_step": WIN,
                   "l2": "each step streams a fresh %d MB window (> nothing cached from the previous step's rows); "
                         "FD state is meant to stay L2-resident" % (WIN * D * 4 // 1000000),
                   "query_level": main_res["level"]},
        "clocks": main_res["clocks"], "multi_stream": multi, "get_ms_per_step": main_res["get_ms"],
        "cpu_batched_baseline": cpu_b, "sharded": sharded,
        "fit_ms_per_step": main_res["fit_ms"], "assign_ms_per_step": main_res["assign_ms"] if args.phase_times else None, "e2e": e2e, "gpu_launches": main_res["launches"], "roofline": roofline,
        "cpu_baseline": cpu, "cov_error": cov,
        "flushes_per_row": sum(o["n_flush"] for o in st) / float(n_steps * WIN),
    }
    if kernels:
        line["kernels"] = kernels
    if other is not None:
        line["reference_schedule"] = {"value": other["value"], "unit": UNIT, "dtype": "f64",
                                      "e2e_value": other.get("e2e_value"),
                                      "ms_per_step": other["ms_total"] / min(2, args.steps),
                                      "k_swfd_rows_share_of_step": (other["k_ms"] / other["ms_total"]) if other.get("k_ms") else None,
                                      "note": "bit-faithful flush/dump schedule of the reference on the same stream "
                                              "(strict parity path, tests/test_swfd_gpu.py)"}
    if cpu and cpu_b:
        # like-for-like ratios (same schedule on both sides); the driver computes its own from the two arms
        line["speedup_same_schedule"] = {
            "strict": (other["value"] / world / cpu["value"]) if other is not None else None,
            "strict_e2e": (other["e2e_value"] / world / cpu["value"]) if other is not None and other.get("e2e_value") else None,
            "batched": main_res["value"] / world / cpu_b["value"],
            "note": "per-GPU rows/s of a schedule / CPU rows/s of the SAME schedule (strict: reference flush schedule vs "
                    "oracle/swfd_oracle.py; batched: vs oracle/swfd_batched_oracle.py)"}
    if args.schedule == "reference":
        ph_tot = {}
        for p in main_res["prof"]:
            for k, v in p.items():
                ph_tot[k] = ph_tot.get(k, 0) + v
        line["phase_cycles_share"] = {k: v / max(1, sum(x for kk, x in ph_tot.items() if kk != "rows"))
                                      for k, v in ph_tot.items() if k != "rows"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    global D, ELL, WIN, R_BOUND, METRIC
    args = parse()
    if args.config == 3:
        # BASELINE.json configs[3]: the shape where K2 / K4 are tensor-bound; same step definition, device legs only
        D, ELL, WIN, R_BOUND = 4096, 256, 100000, 128.0
        METRIC = "stream rows/sec (SWFD update+assign) at d=4096,l=256 (configs[3]); cov-error vs reference"
        args.no_refsched = args.no_sharded = True
        args.streams = ""
        args.no_cpu = args.no_cpu or args.cpu_rows == 0
        if "--steps" not in sys.argv:
            args.steps = 2
        if "--warmup" not in sys.argv:
            args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()

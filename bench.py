#!/usr/bin/env python
"""bench.py -- stream rows/sec of the SWFD update + assign hot path.

Workload (BASELINE.json configs[1]): d=1024, l=128, window N=10k synthetic
stream (seeded low-rank + noise model of swfd/generate_synthetic.py).  A STEP is
one window of 10 000 rows: SeqBasedSWFD.fit of the rows (reference flush / dump
schedule, float64), one SeqBasedSWFD.get() at the window boundary, and the fused
cosine+threshold+argmax assignment of the window's rows against 150 centroids.

  python bench.py [--gpus N] [--steps K] [--warmup W]          # B200 arm
  python bench.py --impl reference ...                          # CPU arm (oracle port)

N > 1 (torchrun): every rank runs an independent stream (the path shards by
independent streams/windows; no data-path collective), value = rows of all
ranks / max-over-ranks time, scaling "weak".
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
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return float(j["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


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
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


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
    cores = host_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(timed) / len(timed),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "SWFD d=1024 l=128 N=10000 low-rank+noise stream + cosine assign k=150 (configs[1])",
                   "schedule": "reference", "levels": sk.logR},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d timed steps of %d rows after %d warm-up rows; one get() of %.3fs amortised over %d rows"
                                   % (args.steps, rows_per_step, warm_fill + args.warmup * rows_per_step, tq, WIN)},
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

    # ---------------- device-resident arm (`value`) -------------------------
    Ad = torch.as_tensor(A).to(dev)                        # inputs resident in HBM before timing
    sk = SeqBasedSWFD(WIN, R, D, ELL, device=local_rank, cluster=args.cluster)
    labels_keep = None

    def step_dev(i):
        blk = Ad[i * WIN:(i + 1) * WIN]
        sk.fit(blk)
        out = sk.get()
        lab = mo.cosine_assign(blk, Cd, TAU, device=dev)
        return out, lab

    for i in range(args.warmup):
        step_dev(i)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    sk.kernel_timing(enable=1)
    l0 = lib.mused_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.warmup, n_steps):
        out, labels_keep = step_dev(i)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = int(lib.mused_launch_count() - l0)
    k_ms, k_launches, k_rows = sk.kernel_timing(enable=0)
    clocks = sampler.stop() if rank == 0 else None
    value = world * args.steps * WIN / (ms_total * 1e-3)
    st = sk.debug_state()
    n_inst = len(st)
    prof = sk.profile()
    last_level = sk.last_level
    sk.close()
    del Ad

    # ---------------- end-to-end arm (`e2e`): host buffers through the public API
    e2e = None
    if not args.no_e2e:
        Ah = torch.as_tensor(A).pin_memory()
        An = Ah.numpy()
        sk2 = SeqBasedSWFD(WIN, R, D, ELL, device=local_rank, cluster=args.cluster)
        lab_host = torch.empty(WIN, dtype=torch.int32).pin_memory()

        def step_e2e(i):
            blk = An[i * WIN:(i + 1) * WIN]
            sk2.fit(blk)                                   # H2D inside the C-ABI call
            out = sk2.get()                                # sketch / sigma^2 / Vt D2H inside
            xd = Ah[i * WIN:(i + 1) * WIN].to(dev, non_blocking=True)
            lab = mo.cosine_assign(xd, Cd, TAU, device=dev)
            lab_host.copy_(lab, non_blocking=True)
            torch.cuda.synchronize()
            return out

        for i in range(args.warmup):
            step_e2e(i)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        for i in range(args.warmup, n_steps):
            step_e2e(i)
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
        e2e = {"value": world * args.steps * WIN / (ms_e2e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": 2 * WIN * D * 4,      # rows for fit + rows for assign
               "d2h_bytes_per_step": 2 * ELL * D * 8 + ELL * 8 + 8 + WIN * 4,
               "wall_s": time.perf_counter() - t0}
        sk2.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel ------------------------
    S = n_inst
    bytes_per_row = 8 * D + S * 20 * D                     # append 8d + S*20d per row
    peak, how = measured_peaks()
    roofline = None
    if k_launches:
        rows_per_launch = k_rows / k_launches
        dur_s = (k_ms / k_launches) * 1e-3
        achieved = bytes_per_row * rows_per_launch / dur_s / 1e9
        roofline = {"bound": "hbm", "kernel": "k_swfd_rows", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": None, "peak_source": how,
                    "algorithmic_bytes_per_row": bytes_per_row, "rows_per_launch": rows_per_launch,
                    "avg_launch_ms": k_ms / k_launches, "share_of_step": k_ms / ms_total,
                    "note": "reference flush schedule is a per-row dependency chain (latency/FP64-bound), see DESIGN.md"}

    cpu = None
    if not args.no_cpu:
        rows = args.cpu_rows or 256
        v, dt = cpu_port_rate(A, R, Cn, warm_rows=256, rows=rows)
        cpu = {"value": v, "unit": UNIT, "cores": host_threads(), "kind": "port",
               "sample": "rows 256..%d of the same stream (after 256 untimed warm-up rows) + one get() + assign, %.1fs"
                         % (256 + rows, dt)}

    ph_tot = {}
    for p in prof:
        for k, v in p.items():
            ph_tot[k] = ph_tot.get(k, 0) + v
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "SWFD d=1024 l=128 N=10000 low-rank+noise stream + cosine assign k=150 (configs[1])",
                   "schedule": "reference", "levels": S // 2, "fd_instances": S, "rows_per_step": WIN,
                   "l2": "each step streams a fresh 41 MB window; FD state (~%.0f MB) is meant to stay L2-resident"
                         % (S * (2 * 2 * ELL * D * 8 + 3 * (2 * ELL + 2) ** 2 * 8) / 1e6),
                   "query_level": last_level},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
        "flushes_per_row": sum(o["n_flush"] for o in st) / float(n_steps * WIN),
        "phase_cycles_share": {k: v / max(1, sum(x for kk, x in ph_tot.items() if kk != "rows"))
                               for k, v in ph_tot.items() if k != "rows"},
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()

"""Independent streams batched on one GPU: K SeqBasedSWFD engines (configs[1] shape), each on its own
CUDA stream, fed window by window; aggregate rows/s against one engine alone."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth

d, ell, N = 1024, 128, 10000
nwin = 3
for K in [int(a) for a in sys.argv[1:]] or [1, 2, 4]:
    engines = [SeqBasedSWFD(N=N, R=64.0, d=d, sketch_dim=ell, schedule="batched") for _ in range(K)]
    streams = [torch.cuda.Stream() for _ in range(K)]
    data = [torch.as_tensor(synth.lowrank_stream((nwin + 1) * N, d, seed=k, scale=2.0)).cuda() for k in range(K)]
    for k in range(K):                      # warm-up window
        with torch.cuda.stream(streams[k]):
            engines[k].fit(data[k][:N])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for w in range(1, nwin + 1):
        # interleave the K engines chunk-block by chunk-block so that their kernels overlap
        for lo in range(w * N, (w + 1) * N, 1280):
            for k in range(K):
                with torch.cuda.stream(streams[k]):
                    engines[k].fit(data[k][lo:lo + 1280])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("K=%d engines: %.0f rows/s aggregate (%.0f per engine)" % (K, K * nwin * N / dt, nwin * N / dt), flush=True)
    for e in engines:
        e.close()

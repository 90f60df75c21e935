#!/bin/bash
# Cycle accounting of one K3 warp (build with -DMUSED_EIG_PROF, run 3 chunks, restore the normal build).
set -e
cd "$(dirname "$0")/.."
MUSED_NVCC_EXTRA=-DMUSED_EIG_PROF python -m mused_b200.build --force > /dev/null
python - <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from mused_b200.swfd import SeqBasedSWFD
from mused_b200 import synth
A = synth.lowrank_stream(10000 + 3 * 128, 1024, seed=0)
sk = SeqBasedSWFD(N=10000, R=64.0, d=1024, sketch_dim=128, schedule="batched")
X = torch.as_tensor(A).cuda()
sk.fit(X[:9984 - 128]); torch.cuda.synchronize()
print("---- one chunk ----", flush=True)
sk.fit(X[9984 - 128:9984]); torch.cuda.synchronize()
PY

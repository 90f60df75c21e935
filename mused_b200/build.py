"""Build libmused_b200.so (hand-written sm_100a CUDA behind a C ABI) in-tree.

``python -m mused_b200.build`` or ``__graft_entry__.build()``.  nvcc
cross-compiles for sm_100a without a GPU.  The .so is git-ignored but travels
with the tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmused_b200.so")
STAMP = os.path.join(HERE, ".libmused_b200.stamp")
SOURCES = ["swfd_engine.cu", "bswfd.cu", "ops.cu", "svd_reduce.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-extended-lambda", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
    "-diag-suppress", "70",
] + os.environ.get("MUSED_NVCC_EXTRA", "").split()      # e.g. -DMUSED_EIG_PROF (cycle accounting of K3)


def _nccl_include():
    """nccl.h of the NCCL that torch bundles (types and prototypes only; the entry points are dlopen'ed at run time)."""
    import importlib.util
    for mod in ("nvidia.nccl", "torch"):
        try:
            spec = importlib.util.find_spec(mod)
        except Exception:
            spec = None
        if spec is None or not spec.submodule_search_locations:
            continue
        base = list(spec.submodule_search_locations)[0]
        for cand in (os.path.join(base, "include"), os.path.join(base, "..", "nvidia", "nccl", "include")):
            if os.path.exists(os.path.join(cand, "nccl.h")):
                return os.path.abspath(cand)
    for cand in ("/usr/include", "/usr/local/cuda/include"):
        if os.path.exists(os.path.join(cand, "nccl.h")):
            return cand
    raise RuntimeError("nccl.h not found (torch's nvidia/nccl/include or /usr/include)")


def _digest():
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for fn in sorted(os.listdir(root)):
            if fn.endswith((".cu", ".cuh", ".h")):
                with open(os.path.join(root, fn), "rb") as f:
                    h.update(fn.encode())
                    h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as f:
            if f.read().strip() == dig:
                return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libmused_b200.so")
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    objs = []
    procs = []
    for s in srcs:
        o = os.path.splitext(s)[0] + ".o"
        objs.append(o)
        cmd = [nvcc, *NVCC_FLAGS, "-I", _nccl_include(), "-c", s, "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), flush=True)
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s" % s)
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB, *objs, "-lcudart", "-lcuda", "-ldl"]
    subprocess.run(cmd, check=True)
    with open(STAMP, "w") as f:
        f.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))

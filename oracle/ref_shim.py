"""Import the UNMODIFIED reference from /root/reference (build container only).

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so
nothing that runs there (``-m gpu`` tests, smoke(), bench.py) may call this;
it is used by ``oracle/gen_golden.py`` and by the CPU tests that re-pin the
restatement when the reference happens to be present.

The reference imports a few packages that are absent here and unused on the
SWFDMC path (swfd/fastfdwithdump.py:4 fbpca; matrix_operations.py:11-12
hdbscan, ot; main.py:11 incdbscan; output_generation.py:2 matplotlib): empty
stub modules are injected for exactly those names.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("MUSED_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "swfd"))


def _stub(name, **attrs):
    if name in sys.modules:
        return
    try:
        importlib.import_module(name)
        return
    except Exception:
        pass
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m


def load():
    """Returns (swfd_pkg, matrix_operations, metrics_evaluation)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    _stub("fbpca", pca=None, eigenn=None)
    _stub("hdbscan")
    _stub("ot")
    _stub("incdbscan", IncrementalDBSCAN=None)
    _stub("matplotlib")
    _stub("matplotlib.pyplot")
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    swfd = importlib.import_module("swfd")
    mo = importlib.import_module("matrix_operations")
    me = importlib.import_module("metrics_evaluation")
    return swfd, mo, me

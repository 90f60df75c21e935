"""CPU: the C-ABI library builds, loads and exports every symbol the header
declares; without a GPU the product path fails loudly (no CPU fallback)."""
import os
import re

import pytest
import torch

from mused_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_and_exports_header_symbols():
    build.build()
    lib = _lib.load()
    hdr = open(os.path.join(ROOT, "include", "mused_b200.h")).read()
    declared = set(re.findall(r"\b(mused_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.mused_version() >= 100


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "mused_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), fn
                assert "scipy.linalg" not in src and "linalg.svd" not in src, fn


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly():
    from mused_b200.swfd import SeqBasedSWFD, FrequentDirectionsWithDump
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        SeqBasedSWFD(100, 20.0, 64, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        FrequentDirectionsWithDump(64, 8, 10.0)


def test_argument_errors_are_value_errors():
    lib = _lib.load()
    if lib.mused_device_count() <= 0:
        pytest.skip("argument validation happens after the device check")
    from mused_b200.swfd import SeqBasedSWFD
    with pytest.raises(ValueError):
        SeqBasedSWFD(100, 20.0, 8, 8)       # d must exceed sketch_dim

"""The C-ABI library loads without a GPU and exports every symbol include/b200lb.h
declares; creating a lattice without a device fails loudly (no CPU fallback)."""

import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200lb.h")).read()
    return sorted(set(re.findall(r"\b(b200lb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    import ludwig_b200 as lb
    L = lb.load()
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), "libb200lb.so does not export %s" % s
    assert sorted(lb.ABI_SYMBOLS) == syms


def test_no_cpu_fallback():
    import torch
    import ludwig_b200 as lb
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lb.B200LBError, match="no CUDA device"):
        lb.Lattice(19, (4, 4, 4))


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ludwig_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".c")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in text.replace("no oracle", ""), fn

"""CPU test: the C-ABI library loads and exports every symbol include/fempar_b200.h declares; without a GPU
the product fails loudly instead of falling back."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "fempar_b200.h")).read()
    return sorted(set(re.findall(r"\b(fb2_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    import fempar_b200
    L = ctypes.CDLL(fempar_b200._lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 45
    for n in names:
        assert hasattr(L, n), n
    assert set(fempar_b200._lib.EXPORTED_SYMBOLS) <= set(names)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import fempar_b200
    with pytest.raises(fempar_b200.Fb2Error):
        fempar_b200.init(0)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "fempar_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                assert "oracle" not in open(os.path.join(dirpath, f)).read().lower(), f

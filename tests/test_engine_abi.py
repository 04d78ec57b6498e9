"""The C-ABI library loads and exports every symbol include/mgengine.h declares (no GPU needed)."""
import ctypes
import os
import re

import pytest

from evostencils_b200 import engine

HEADER = os.path.join(os.path.dirname(__file__), "..", "include", "mgengine.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mg_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(engine.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol():
    if not os.path.exists(engine.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(engine.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(engine.EngineError, match="no CUDA device|CUDA"):
        engine.Workspace(2, 2, 4, 1)

"""CPU-only: the C-ABI library loads and exports every symbol include/*.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(lime_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from lime_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared("lime_cuda.h") + _declared("lime_host.h")
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), n
    # and the binding lists exactly those
    assert sorted(_lib.SIGNATURES) == sorted(names)
    _lib.load()


def test_no_cuda_device_is_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        return
    import pytest
    from lime_b200.api import LimeContext, LimeError
    with pytest.raises(LimeError):
        LimeContext(0)


def test_product_does_not_import_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "lime_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in src.lower() or f == "_lib.py" and False, os.path.join(dirpath, f)

"""CPU tier: the C-ABI library loads, exports every symbol include/gpqhe_b200.h declares,
and fails loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from gpqhe_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "gpqhe_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gq_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    syms = declared_symbols()
    assert len(syms) >= 40
    assert set(syms) == set(capi.SIGNATURES), set(syms) ^ set(capi.SIGNATURES)


def test_real_library_exports_every_symbol(real_lib):
    lib = ctypes.CDLL(real_lib)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} missing from libgpqhe_b200.so"
    assert capi.load(real_lib).gq_abi_version() == 1


def test_no_cpu_fallback(real_lib):
    """Without a CUDA device context creation must raise, not fall back."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    import gpqhe_b200 as g
    with pytest.raises(g.GpqheError):
        g.HeContext(8, 1 << 120, 4, 1 << 40, lib_path=real_lib)


def test_missing_library_is_loud(tmp_path):
    with pytest.raises(capi.GpqheError):
        capi.load(str(tmp_path / "nope.so"))

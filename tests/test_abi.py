"""CPU: libmm25.so builds for sm_100a, loads, and exports every symbol include/mm25.h declares;
the host wrappers refuse CPU tensors (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "mm25.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mm25_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(lib_built):
    from multimodal_2025_b200 import _lib
    h = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(h, s), "libmm25.so lacks %s" % s
        assert s in _lib.SIGNATURES, "ctypes signature missing for %s" % s
    assert set(_lib.SIGNATURES) == set(syms)
    assert b"sm_100a" in h.mm25_version.__call__() if False else True


def test_version_string(lib_built):
    from multimodal_2025_b200 import _lib
    assert b"sm_100a" in _lib.lib().mm25_version()


def test_cpu_tensors_raise(lib_built):
    from multimodal_2025_b200.losses import ClipLoss
    m = ClipLoss()
    x = torch.randn(8, 16)
    with pytest.raises((NotImplementedError, RuntimeError)):
        m(x, x, torch.tensor(2.0))


def test_module_surface_matches_reference(lib_built):
    import inspect
    from multimodal_2025_b200.losses import ClipLoss
    m = ClipLoss()
    assert "smoothing" in m.state_dict()           # clipLoss.py:22
    assert abs(float(m.state_dict()["smoothing"]) - 0.1) < 1e-7
    assert list(inspect.signature(m.forward).parameters) == [
        "image_features", "text_features", "logits_scale", "normalize"]
    assert list(inspect.signature(m.forward_sampled).parameters) == [
        "image_features", "text_features", "logits_scale"]

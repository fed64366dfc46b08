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
    h.mm25_version.restype = ctypes.c_char_p
    assert b"sm_100a" in h.mm25_version()


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


def _our_surface():
    """qualified name -> callable of the drop-in modules, keyed like tests/golden/surface.json."""
    from multimodal_2025_b200.losses import alignCLIPLoss, align_clip_loss, clipLoss, cloobLoss
    from multimodal_2025_b200.losses import vclipLoss
    from multimodal_2025_b200.losses.distributions import ProjectedNormal, VonMisesFisher
    from multimodal_2025_b200.metrics import entropy, retrieval, topk
    S = {
        "ClipLoss.__init__": clipLoss.ClipLoss.__init__,
        "ClipLoss.forward": clipLoss.ClipLoss.forward,
        "ClipLoss.forward_sampled": clipLoss.ClipLoss.forward_sampled,
        "CLOOBLoss.__init__": cloobLoss.CLOOBLoss.__init__,
        "CLOOBLoss.forward": cloobLoss.CLOOBLoss.forward,
        "VClipLoss.__init__": vclipLoss.VClipLoss.__init__,
        "VClipLoss.forward": vclipLoss.VClipLoss.forward,
        "ClipInModalityLoss.__init__": align_clip_loss.ClipInModalityLoss.__init__,
        "ClipInModalityLoss.forward": align_clip_loss.ClipInModalityLoss.forward,
        "align.ClipLoss.__init__": align_clip_loss.ClipLoss.__init__,
        "align.ClipLoss.forward": align_clip_loss.ClipLoss.forward,
        "AlignCLIPLoss.__init__": alignCLIPLoss.AlignCLIPLoss.__init__,
        "AlignCLIPLoss.forward": alignCLIPLoss.AlignCLIPLoss.forward,
        "VonMisesFisher.__init__": VonMisesFisher.__init__,
        "ProjectedNormal.__init__": ProjectedNormal.__init__,
        "TopKMetric.__init__": topk.TopKMetric.__init__,
        "TopKMetric.compute": topk.TopKMetric.compute,
        "topk.top_k_similarities": topk.top_k_similarities,
        "topk.top_k_score": topk.top_k_score,
        "EntropyMetric.__init__": entropy.EntropyMetric.__init__,
        "EntropyMetric.compute": entropy.EntropyMetric.compute,
        "entropy.average_kth_neighbor_angle": entropy.average_kth_neighbor_angle,
        "train.get_clip_metrics": retrieval.get_clip_metrics,
        "retrieval.recall_at_k": retrieval.recall_at_k,
    }
    for f in ("cloob_loss", "infoloob_loss", "hopfield_retrieval", "hopfield"):
        S["cloobLoss." + f] = getattr(cloobLoss, f)
    for f in ("power_spherical_mean", "power_spherical_mean2", "power_spherical_mean_normalized",
              "vmf_mean_normalized", "get_distribution_mean"):
        S["vclipLoss." + f] = getattr(vclipLoss, f)
    for m in ("rsample", "sample", "log_prob", "entropy", "expand"):
        S["VonMisesFisher." + m] = getattr(VonMisesFisher, m)
    for m in ("rsample", "sample", "log_prob", "entropy"):
        S["ProjectedNormal." + m] = getattr(ProjectedNormal, m)
    return S


def test_every_dropin_signature_matches_the_recorded_reference_surface(lib_built):
    """same parameter names, order and defaults as the reference (recorded from the
    reference itself by tests/golden/make_surface.py).  The drop-in may only ADD trailing
    parameters that have defaults (e.g. `device=None`, `encode_dataset=None`)."""
    import inspect
    import json
    import os
    from conftest import GOLDEN
    ref = json.load(open(os.path.join(GOLDEN, "surface.json")))
    ours = _our_surface()
    assert set(ref) <= set(ours), sorted(set(ref) - set(ours))
    for name, want in ref.items():
        if name.endswith("__init__") and ours[name] is object.__init__:
            assert want == [], name          # plain class without a constructor on both sides
            continue
        ps = [p for p in inspect.signature(ours[name]).parameters.values() if p.name != "self"]
        got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
               for p in ps]
        if want == [] and [g[0] for g in got] == ["args", "kwargs"]:
            continue                         # object.__init__ seen through a class
        assert [g[0] for g in got[:len(want)]] == [w[0] for w in want], (name, got, want)
        for g, w in zip(got, want):
            if w[1] is None:
                # a required reference parameter may have gained a default (device=None)
                continue
            assert g[1] is not None and _same_default(g[1], w[1]), (name, g, w)
        for extra in got[len(want):]:
            assert extra[1] is not None, (name, "extra required parameter", extra)


def _same_default(a, b):
    if a == b:
        return True
    try:
        return float(a) == float(b)
    except ValueError:
        return False


def _header_prototypes():
    """name -> (return C type, [parameter C types]) parsed from include/mm25.h."""
    src = open(os.path.join(ROOT, "include", "mm25.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    out = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(mm25_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, re.S):
        ret, name, params = m.group(1).strip(), m.group(2), m.group(3).strip()
        ptypes = []
        if params and params != "void":
            for p in params.split(","):
                p = " ".join(p.split())
                if "*" in p:
                    ptypes.append("ptr")
                else:
                    ptypes.append(" ".join(p.split(" ")[:-1]))     # drop the parameter name
        out[name] = (ret, ptypes)
    return out


def test_ctypes_signatures_agree_with_the_header(lib_built):
    """every binding in _lib.SIGNATURES has the header's arity and, per position, a ctypes type of
    the header's C type (pointer / int64_t / int / float / double / size_t / uint64_t) — a
    mismatch here would corrupt arguments silently on the GPU box."""
    from multimodal_2025_b200 import _lib
    cmap = {"int64_t": ctypes.c_int64, "int": ctypes.c_int, "float": ctypes.c_float,
            "double": ctypes.c_double, "size_t": ctypes.c_size_t, "uint64_t": ctypes.c_uint64,
            "ptr": ctypes.c_void_p, "long long": ctypes.c_longlong}
    rmap = {"int": ctypes.c_int, "size_t": ctypes.c_size_t, "int64_t": ctypes.c_int64,
            "void": None, "const char*": ctypes.c_char_p, "const char *": ctypes.c_char_p,
            "long long": ctypes.c_longlong}
    protos = _header_prototypes()
    assert set(protos) == set(_lib.SIGNATURES)
    for name, (ret, ptypes) in protos.items():
        res, args = _lib.SIGNATURES[name]
        assert rmap[ret] == res, (name, ret, res)
        assert len(args) == len(ptypes), (name, len(args), len(ptypes))
        for i, (a, p) in enumerate(zip(args, ptypes)):
            assert cmap[p] == a, (name, i, p, a)

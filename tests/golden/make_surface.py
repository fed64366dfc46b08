"""Record the call signatures of the reference's public surface on this path by
IMPORTING THE REFERENCE and inspecting it.  Build container only:
    python tests/golden/make_surface.py
Writes tests/golden/surface.json = {qualified name: [[param, default-repr or null], ...]}.
tests/test_abi.py compares the drop-in modules with it (the GPU box has no /root/reference).

Shims as in make_golden*.py: `power_spherical`, `traitlets` are absent here; AlignCLIP's loss
module is loaded by file path; `get_clip_metrics` / `recall_at_k` are taken from the `ast` of
their own source files (whose imports need wandb / align_clip).
"""
import ast
import importlib.util
import inspect
import json
import os
import sys
import types
import typing

REF = os.environ.get("MM25_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)

shim = types.ModuleType("power_spherical")
shim.PowerSpherical = type("PowerSpherical", (), {})
shim.HypersphericalUniform = type("HypersphericalUniform", (), {})
sys.modules["power_spherical"] = shim
tl = types.ModuleType("traitlets")
tl.List = typing.List
sys.modules["traitlets"] = tl


def by_path(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def sig(fn):
    out = []
    for p in inspect.signature(fn).parameters.values():
        if p.name == "self":
            continue
        out.append([p.name, None if p.default is inspect.Parameter.empty else repr(p.default)])
    return out


def ast_sig(rel, name):
    tree = ast.parse(open(os.path.join(REF, rel)).read())
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name)
    args = node.args
    names = [a.arg for a in args.args]
    defaults = [None] * (len(names) - len(args.defaults)) + [ast.unparse(d) for d in args.defaults]
    return [[n, d] for n, d in zip(names, defaults)]


def main():
    import losses.clipLoss as clip
    import losses.cloobLoss as cloob
    import losses.vclipLoss as vclip
    import losses.distributions.VonMisesFisher as vmf
    import losses.distributions.ProjectedNormal as pn
    import metrics.topk as topk
    import metrics.entropy as entropy
    align = by_path("ref_alignloss", "AlignCLIP/align_clip/loss.py")
    # losses/alignCLIPLoss.py imports the AlignCLIP package (needs ftfy): read it by ast
    S = {}
    S["ClipLoss.__init__"] = sig(clip.ClipLoss.__init__)
    S["ClipLoss.forward"] = sig(clip.ClipLoss.forward)
    S["ClipLoss.forward_sampled"] = sig(clip.ClipLoss.forward_sampled)
    S["CLOOBLoss.__init__"] = sig(cloob.CLOOBLoss.__init__)
    S["CLOOBLoss.forward"] = sig(cloob.CLOOBLoss.forward)
    for f in ("cloob_loss", "infoloob_loss", "hopfield_retrieval", "hopfield"):
        S["cloobLoss." + f] = sig(getattr(cloob, f))
    S["VClipLoss.__init__"] = sig(vclip.VClipLoss.__init__)
    S["VClipLoss.forward"] = sig(vclip.VClipLoss.forward)
    for f in ("power_spherical_mean", "power_spherical_mean2", "power_spherical_mean_normalized",
              "vmf_mean_normalized", "get_distribution_mean"):
        S["vclipLoss." + f] = sig(getattr(vclip, f))
    S["ClipInModalityLoss.__init__"] = sig(align.ClipInModalityLoss.__init__)
    S["ClipInModalityLoss.forward"] = sig(align.ClipInModalityLoss.forward)
    S["align.ClipLoss.__init__"] = sig(align.ClipLoss.__init__)
    S["align.ClipLoss.forward"] = sig(align.ClipLoss.forward)
    tree = ast.parse(open(os.path.join(REF, "losses/alignCLIPLoss.py")).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "AlignCLIPLoss")
    for fn in cls.body:
        if isinstance(fn, ast.FunctionDef) and fn.name in ("__init__", "forward"):
            names = [a.arg for a in fn.args.args if a.arg != "self"]
            dfl = [None] * (len(names) - len(fn.args.defaults)) + [ast.unparse(d)
                                                                   for d in fn.args.defaults]
            S["AlignCLIPLoss." + fn.name] = [[n, d] for n, d in zip(names, dfl)]
    S["VonMisesFisher.__init__"] = sig(vmf.VonMisesFisher.__init__)
    for m in ("rsample", "sample", "log_prob", "entropy", "expand"):
        S["VonMisesFisher." + m] = sig(getattr(vmf.VonMisesFisher, m))
    S["ProjectedNormal.__init__"] = sig(pn.ProjectedNormal.__init__)
    for m in ("rsample", "sample", "log_prob", "entropy"):
        if hasattr(pn.ProjectedNormal, m):
            S["ProjectedNormal." + m] = sig(getattr(pn.ProjectedNormal, m))
    S["TopKMetric.__init__"] = sig(topk.TopKMetric.__init__)
    S["TopKMetric.compute"] = sig(topk.TopKMetric.compute)
    S["topk.top_k_similarities"] = sig(topk.top_k_similarities)
    S["topk.top_k_score"] = sig(topk.top_k_score)
    S["EntropyMetric.__init__"] = sig(entropy.EntropyMetric.__init__)
    S["EntropyMetric.compute"] = sig(entropy.EntropyMetric.compute)
    S["entropy.average_kth_neighbor_angle"] = sig(entropy.average_kth_neighbor_angle)
    S["train.get_clip_metrics"] = ast_sig("AlignCLIP/main/train.py", "get_clip_metrics")
    S["retrieval.recall_at_k"] = ast_sig("AlignCLIP/main/retrieval.py", "recall_at_k")
    with open(os.path.join(OUT, "surface.json"), "w") as f:
        json.dump(S, f, indent=1, sort_keys=True)
    print("recorded", len(S), "signatures")


if __name__ == "__main__":
    main()

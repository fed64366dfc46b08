"""Golden vectors for the retrieval metrics, produced by EXECUTING THE REFERENCE's own functions.

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden_metrics.py
Writes tests/golden/metrics.npz.

How the reference code is reached without its heavy imports (wandb, align_clip, ftfy ...):
  * `AlignCLIP/main/train.py::get_clip_metrics` and `AlignCLIP/main/retrieval.py::recall_at_k`
    are compiled from their own source text (the `ast` node of that one function) into a
    namespace that provides `torch`, `np`, `List`; nothing is copied into this repo.
  * `recall_at_k` calls `encode_dataset` (encoders: out of scope) and `.cuda()` (no GPU here):
    the namespace provides an `encode_dataset` that returns the prepared encodings, and
    `Tensor.cuda` is patched to the identity while the function runs.
  * `metrics/entropy.py` is imported as a module with a 2-line `traitlets` stub.
All inputs are checked to hold no score tied with a ranked target (the only case in which
`torch.argsort` gives the reference a defined answer for that target).
"""
import ast
import os
import sys
import types
import typing

import numpy as np
import torch
import torch.nn.functional as F

REF = os.environ.get("MM25_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def function_from_source(rel, name, namespace):
    src = open(os.path.join(REF, rel)).read()
    tree = ast.parse(src)
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name)
    mod = ast.Module(body=[node], type_ignores=[])
    ns = dict(namespace)
    exec(compile(mod, os.path.join(REF, rel), "exec"), ns)
    return ns[name], ns


def tie_free(m, row_targets):
    """True when no other entry of row q equals m[q, row_targets[q]] — the position of a target
    in a sort depends only on ties WITH the target.  row_targets: [Q] or [Q, c]."""
    t = row_targets.view(m.shape[0], -1)
    for c in range(t.shape[1]):
        tv = m.gather(1, t[:, c:c + 1])
        if bool(((m == tv).sum(dim=1) != 1).any()):
            return False
    return True


def main():
    torch.set_num_threads(4)
    res = {}

    # ---- get_clip_metrics (train.py:338-356) ------------------------------------------------
    gcm, _ = function_from_source("AlignCLIP/main/train.py", "get_clip_metrics",
                                  {"torch": torch, "np": np})
    for tag, (n, d, seed, noise, ls) in {"n300_d64": (300, 64, 11, 0.9, 14.285714),
                                         "n97_d32": (97, 32, 12, 1.5, 100.0),
                                         "n1030_d48": (1030, 48, 13, 1.2, 1.0)}.items():
        g = torch.Generator().manual_seed(seed)
        img = F.normalize(torch.randn(n, d, generator=g), dim=-1)
        txt = F.normalize(img + noise * torch.randn(n, d, generator=g) / d ** 0.5, dim=-1)
        logits = ls * img @ txt.t()
        assert tie_free(logits, torch.arange(n)) and tie_free(logits.t(), torch.arange(n)), tag
        m = gcm(img, txt, ls)
        res[f"gcm_{tag}/img"] = img.numpy()
        res[f"gcm_{tag}/txt"] = txt.numpy()
        res[f"gcm_{tag}/logit_scale"] = np.float64(ls)
        for k, v in m.items():
            res[f"gcm_{tag}/{k}"] = np.float64(v)
        # the per-item ranks the reference derives on the way (train.py:349-350)
        for name, lg in (("image_to_text", logits), ("text_to_image", logits.t())):
            ranking = torch.argsort(lg, descending=True)
            res[f"gcm_{tag}/{name}_preds"] = torch.where(
                ranking == torch.arange(n).view(-1, 1))[1].numpy().astype(np.int64)

    # ---- recall_at_k (retrieval.py:55-102) --------------------------------------------------
    holder = {}

    def fake_encode_dataset(clip, dataset, device, batch_size=16):
        return holder["enc"]

    dutils = types.SimpleNamespace(Dataset=object)
    rak, _ = function_from_source("AlignCLIP/main/retrieval.py", "recall_at_k",
                                  {"torch": torch, "List": typing.List, "dutils": dutils,
                                   "encode_dataset": fake_encode_dataset, "print": lambda *a: 0})
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for tag, (num_im, cpi, d, seed, noise) in {"im120_c5_d40": (120, 5, 40, 21, 1.0),
                                                   "im64_c1_d24": (64, 1, 24, 22, 2.0),
                                                   "im257_c3_d32": (257, 3, 32, 23, 1.4)}.items():
            g = torch.Generator().manual_seed(seed)
            im = F.normalize(torch.randn(num_im, d, generator=g), dim=-1)
            tx = im.repeat_interleave(cpi, dim=0) + noise * torch.randn(num_im * cpi, d,
                                                                        generator=g) / d ** 0.5
            tx = F.normalize(tx, dim=-1)
            t2i = torch.arange(num_im).repeat_interleave(cpi)
            i2t = torch.arange(num_im * cpi).view(num_im, cpi)
            assert tie_free(tx @ im.T, t2i) and tie_free((tx @ im.T).T, i2t), tag
            k_vals = [1, 5, 10, 50]
            holder["enc"] = (im, tx, t2i, i2t)
            t2i_recall, i2t_recall = rak(None, None, "cpu", k_vals, 16)
            res[f"rak_{tag}/im"] = im.numpy()
            res[f"rak_{tag}/tx"] = tx.numpy()
            res[f"rak_{tag}/t2i_map"] = t2i.numpy().astype(np.int64)
            res[f"rak_{tag}/i2t_map"] = i2t.numpy().astype(np.int64)
            res[f"rak_{tag}/k_vals"] = np.array(k_vals, dtype=np.int64)
            res[f"rak_{tag}/t2i_recall"] = np.array(t2i_recall, dtype=np.float64)
            res[f"rak_{tag}/i2t_recall"] = np.array(i2t_recall, dtype=np.float64)
    finally:
        torch.Tensor.cuda = orig_cuda

    # ---- entropy proxy (metrics/entropy.py:37-80) -------------------------------------------
    tl = types.ModuleType("traitlets")
    tl.List = typing.List
    sys.modules["traitlets"] = tl
    sys.path.insert(0, REF)
    from metrics import entropy as ref_entropy  # noqa: E402
    for tag, (n, d, seed, k, bs) in {"n500_d32_k5": (500, 32, 31, 5, 1000),
                                     "n333_d16_k1_b100": (333, 16, 32, 1, 100),
                                     "n130_d8_k20_b64": (130, 8, 33, 20, 64)}.items():
        g = torch.Generator().manual_seed(seed)
        emb = torch.randn(n, d, generator=g) * (0.5 + torch.rand(n, 1, generator=g))
        val = ref_entropy.average_kth_neighbor_angle(emb, k, bs)
        res[f"ent_{tag}/emb"] = emb.numpy()
        res[f"ent_{tag}/k"] = np.int64(k)
        res[f"ent_{tag}/batch_size"] = np.int64(bs)
        res[f"ent_{tag}/value"] = np.float64(val)
    pairs = [([torch.tensor([1.0, 0.0, 0.0]), torch.tensor([0.0, 1.0, 0.0])],
              torch.tensor([0.0, 0.0, 1.0])),
             ([torch.tensor([1.0, 1.0, 0.0])], torch.tensor([1.0, 0.2, 0.9]))]
    res["ent_metric/value_k1"] = np.float64(ref_entropy.EntropyMetric(k=1).compute(pairs))
    res["ent_metric/value_k3"] = np.float64(ref_entropy.EntropyMetric(k=3).compute(pairs))

    np.savez_compressed(os.path.join(OUT, "metrics.npz"), **res)
    print("metrics golden vectors written:", len(res), "arrays")


if __name__ == "__main__":
    main()

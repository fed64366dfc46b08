"""Generate golden vectors by RUNNING THE REFERENCE ITSELF (imported from /root/reference).

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden.py
Writes tests/golden/*.npz (inputs + reference outputs + reference gradients, fp32).  The
oracle (oracle/losses_oracle.py) and the CUDA path are both checked against these files.

Shims: `power_spherical` and `traitlets` are absent here, AlignCLIP's package
__init__ needs `ftfy`; the loss modules themselves are loaded unmodified by file path.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

REF = os.environ.get("MM25_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def load_by_path(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def synth(n, d, seed, correlated=False):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = F.normalize(torch.randn(n, d, generator=g), dim=-1)
    if correlated:
        y = F.normalize(x + 0.5 * torch.randn(n, d, generator=g), dim=-1)
    else:
        y = F.normalize(torch.randn(n, d, generator=g), dim=-1)
    return x, y


def run(fn, *tensors):
    leaves = [t.detach().clone().requires_grad_(True) for t in tensors]
    out = fn(*leaves)
    out.sum().backward()
    return out.detach().numpy(), [l.grad.numpy() for l in leaves]


def main():
    torch.set_num_threads(4)
    clip_mod = load_by_path("ref_clipLoss", "losses/clipLoss.py")
    cloob_mod = load_by_path("ref_cloobLoss", "losses/cloobLoss.py")
    align_mod = load_by_path("ref_alignloss", "AlignCLIP/align_clip/loss.py")

    # ---- CLIP forward (clipLoss.py:81-121) --------------------------------------------------
    cases = {}
    for tag, (n, d, seed, corr, ls, eps, raw) in {
        "n256_d512": (256, 512, 0, False, 2.6592, 0.1, False),       # BASELINE configs[0]
        "n256_d512_clamp": (256, 512, 1, True, 5.0, 0.1, False),      # above the 4.6052 clamp
        "n192_d96_eps0": (192, 96, 2, True, 3.2, 0.0, False),         # ragged vs 64/128 tiles
        "n70_d40_raw": (70, 40, 3, False, 2.0, 0.1, True),            # un-normalised inputs
    }.items():
        x, y = synth(n, d, seed, corr)
        if raw:
            g = torch.Generator().manual_seed(seed + 100)
            x = x * (0.5 + torch.rand(n, 1, generator=g) * 3)
            y = y * (0.5 + torch.rand(n, 1, generator=g) * 3)
        lsc = torch.tensor(ls)
        m = clip_mod.ClipLoss(eps)
        loss, (gx, gy, gl) = run(lambda a, b, c: m(a, b, c), x, y, lsc)
        cases[tag] = dict(x=x.numpy(), y=y.numpy(), ls=np.float32(ls), eps=np.float32(eps),
                          loss=loss, gx=gx, gy=gy, gl=gl)
    # normalize=False branch
    x, y = synth(96, 64, 7, True)
    m = clip_mod.ClipLoss(0.1)
    lsc = torch.tensor(2.5)
    loss, (gx, gy, gl) = run(lambda a, b, c: m(a, b, c, normalize=False), x * 1.3, y * 0.8, lsc)
    cases["n96_d64_nonorm"] = dict(x=(x * 1.3).numpy(), y=(y * 0.8).numpy(), ls=np.float32(2.5),
                                   eps=np.float32(0.1), loss=loss, gx=gx, gy=gy, gl=gl)
    np.savez_compressed(os.path.join(OUT, "clip_forward.npz"),
                        **{f"{t}/{k}": v for t, c in cases.items() for k, v in c.items()})

    # ---- CLIP forward_sampled (clipLoss.py:24-79) ------------------------------------------
    g = torch.Generator().manual_seed(11)
    xs = torch.randn(3, 80, 48, generator=g)
    ys = xs + 2.5 * torch.randn(3, 80, 48, generator=g)
    out = {}
    for eps in (0.1, 0.0):
        m = clip_mod.ClipLoss(eps)
        for ls in (2.6592, 4.2):
            w = torch.tensor([1.0, 0.5, 2.0])
            loss, (gx, gy, gl) = run(lambda a, b, c: (m.forward_sampled(a, b, c) * w), xs, ys,
                                     torch.tensor(ls))
            out[f"eps{eps}_ls{ls}"] = dict(loss=loss / w.numpy(), gx=gx, gy=gy, gl=gl)
    np.savez_compressed(os.path.join(OUT, "clip_sampled.npz"), x=xs.numpy(), y=ys.numpy(),
                        w=np.array([1.0, 0.5, 2.0], dtype=np.float32),
                        **{f"{t}/{k}": v for t, c in out.items() for k, v in c.items()})

    # ---- CLOOB (cloobLoss.py:10-61) ---------------------------------------------------------
    out = {}
    for tag, (n, d, seed, corr) in {"n128_d64": (128, 64, 4, True),
                                    "n200_d96": (200, 96, 5, False)}.items():
        x, y = synth(n, d, seed, corr)
        m = cloob_mod.CLOOBLoss(30.0, 15.0, "cpu")
        loss, (gx, gy) = run(lambda a, b: m(a, b), x, y)
        pxx = cloob_mod.hopfield(x, x, 15.0).numpy()
        pxy = cloob_mod.hopfield(y, x, 15.0).numpy()
        out[tag] = dict(x=x.numpy(), y=y.numpy(), loss=loss, gx=gx, gy=gy, pxx=pxx, pxy=pxy)
    np.savez_compressed(os.path.join(OUT, "cloob.npz"),
                        **{f"{t}/{k}": v for t, c in out.items() for k, v in c.items()})

    # ---- AlignCLIP (loss.py) — the branches that run unmodified on CPU ----------------------
    out = {}
    x, y = synth(160, 64, 6, True)
    s = torch.tensor(14.285)
    crit = align_mod.ClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=False)
    loss, (gx, gy, gs) = run(lambda a, b, c: crit(a, b, c), x, y, s)
    out["nosem"] = dict(x=x.numpy(), y=y.numpy(), s=np.float32(14.285), loss=loss, gx=gx, gy=gy,
                        gs=gs)
    crit2 = align_mod.ClipLoss()
    loss, (gx, gy, gs) = run(lambda a, b, c: crit2(a, b, c), x, y, s)
    out["openclip"] = dict(loss=loss, gx=gx, gy=gy, gs=gs)
    np.savez_compressed(os.path.join(OUT, "alignclip.npz"),
                        **{f"{t}/{k}": v for t, c in out.items() for k, v in c.items()})

    # ---- top-k (metrics/topk.py:30-82 and its __main__ KAT :85-106) -------------------------
    tl = types.ModuleType("traitlets")
    import typing
    tl.List = typing.List
    sys.modules["traitlets"] = tl
    sys.path.insert(0, REF)
    from metrics import topk as ref_topk  # noqa: E402
    emb1 = torch.tensor([[1.0, 0.0], [0.9, 0.1], [0.0, 1.0]])
    query = torch.tensor([1.0, 0.0])
    kat = ref_topk.top_k_similarities(emb1, query, k=2)
    score = ref_topk.top_k_score([([torch.tensor([1.0, 0.0]), torch.tensor([0.9, 0.1]),
                                    torch.tensor([0.0, 1.0])], torch.tensor([1.0, 0.0]))], k=1)
    g = torch.Generator().manual_seed(21)
    texts = torch.randn(300, 32, generator=g)
    images = texts[::3] + 0.8 * torch.randn(100, 32, generator=g)
    pairs = [([texts[3 * i], texts[3 * i + 1], texts[3 * i + 2]], images[i]) for i in range(100)]
    res = {}
    for k in (1, 5, 10, 100):
        acc, info = ref_topk.top_k_score(pairs, k)
        idx = np.array([[i for i, _ in ref_topk.top_k_similarities(texts, images[q], k)]
                        for q in range(100)], dtype=np.int64)
        res[f"k{k}/acc"] = np.float64(acc)
        res[f"k{k}/correct"] = np.int64(info["correct"])
        res[f"k{k}/idx"] = idx
    np.savez_compressed(os.path.join(OUT, "topk.npz"), texts=texts.numpy(), images=images.numpy(),
                        kat_idx=np.array([i for i, _ in kat], dtype=np.int64),
                        kat_sim=np.array([s for _, s in kat], dtype=np.float32),
                        kat_score=np.float64(score[0]), **res)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()

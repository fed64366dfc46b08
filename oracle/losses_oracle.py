"""CPU oracle — TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py; never by the product path).

A torch-CPU restatement of the reference's loss algorithms, line by line, with the working
dtype selectable (fp32 = what the reference computes, fp64 = ground truth for error budgets).
Gradients come from autograd on this restatement.  Pinned against the reference itself: the
golden vectors in tests/golden/*.npz were produced by tests/golden/make_golden.py importing the
real modules from /root/reference (losses/clipLoss.py, losses/cloobLoss.py,
AlignCLIP/align_clip/loss.py); tests/test_oracle.py checks this file against them.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------------
# CLIP — reference losses/clipLoss.py
# --------------------------------------------------------------------------------------------
def clip_loss(image_features, text_features, logits_scale, label_smoothing=0.1, normalize=True,
              dtype=torch.float32):
    """clipLoss.py:81-121 (ClipLoss.forward)."""
    if normalize:  # :94-96
        image_features = F.normalize(image_features, dim=-1, p=2)
        text_features = F.normalize(text_features, dim=-1, p=2)
    batch_size = image_features.shape[0]  # :102
    logit_scale = logits_scale.clamp(max=4.6052).exp()  # :105
    logits_per_image = logit_scale.to(dtype) * image_features.to(dtype) @ text_features.t().to(dtype)  # :106
    logits_per_text = logits_per_image.t()  # :107
    labels = torch.arange(batch_size, device=image_features.device)  # :113
    loss_img = F.cross_entropy(logits_per_image, labels, label_smoothing=label_smoothing)  # :116
    loss_txt = F.cross_entropy(logits_per_text, labels, label_smoothing=label_smoothing)  # :117
    return (loss_img + loss_txt) / 2  # :119


def clip_loss_slab(image_features, text_features, logits_scale, row0, rows, label_smoothing=0.1,
                   normalize=True, dtype=torch.float32):
    """The part of `clip_loss` (clipLoss.py:94-119) owned by image rows [row0, row0 + rows): the
    [rows, N] slab of the logits (:106) with the same per-logit arithmetic.  Returns the pieces
    that `clip_loss_from_slabs` combines into exactly the loss of the whole batch:
      img_sum   = sum over the slab's rows of the image-side CE (:116, label smoothing included)
      col_lse   = [N] logsumexp over the slab's rows of every column (partial column statistic)
      col_lin   = [N] sum over the slab's rows of z_ij   (smoothing term of the text-side CE)
      diag      = [rows] the positives z_ii
    bench.py times this as a bounded sample of a batch too large to run whole on the host: a
    slab of r rows is r / N of the N x N work with the per-pair work unchanged."""
    if normalize:
        image_features = F.normalize(image_features, dim=-1, p=2)
        text_features = F.normalize(text_features, dim=-1, p=2)
    logit_scale = logits_scale.clamp(max=4.6052).exp()
    z = logit_scale.to(dtype) * image_features[row0:row0 + rows].to(dtype) @ text_features.t().to(dtype)
    labels = torch.arange(row0, row0 + rows, device=z.device)
    img_sum = F.cross_entropy(z, labels, label_smoothing=label_smoothing, reduction="sum")
    col_lse = torch.logsumexp(z, dim=0)
    col_lin = z.sum(dim=0)
    diag = z[torch.arange(rows), labels]
    return img_sum, col_lse, col_lin, diag


def clip_loss_from_slabs(parts, n, label_smoothing=0.1):
    """Combine the pieces of `clip_loss_slab` over a partition of the rows into clip_loss."""
    img = sum(p[0] for p in parts) / n
    col_lse = torch.logsumexp(torch.stack([p[1] for p in parts]), dim=0)
    col_lin = torch.stack([p[2] for p in parts]).sum(dim=0)
    diag = torch.cat([p[3] for p in parts])
    eps = label_smoothing
    # text-side CE with label smoothing: mean_j [ lse_j - (1 - eps) z_jj - eps / n * sum_i z_ij ]
    txt = (col_lse - (1 - eps) * diag - eps / n * col_lin).mean()
    return (img + txt) / 2


def clip_loss_sampled(image_features, text_features, logits_scale, label_smoothing=0.1):
    """clipLoss.py:24-79 (ClipLoss.forward_sampled): [S,B,E] x2 -> [S]."""
    B1, B2, E = image_features.shape
    image_features = F.normalize(image_features, dim=-1, p=2)  # :28
    text_features = F.normalize(text_features, dim=-1, p=2)  # :29
    logit_scale = logits_scale.clamp(max=3.912).exp()  # :35
    logits_per_image = logit_scale * image_features @ text_features.permute(0, 2, 1)  # :38
    logits_per_text = logits_per_image.permute(0, 2, 1)  # :39
    labels = torch.arange(B2, device=image_features.device).expand(B1, B2)  # :45
    li = logits_per_image.reshape(B1 * B2, B2)  # :50
    lt = logits_per_text.reshape(B1 * B2, B2)  # :51
    lab = labels.reshape(B1 * B2)  # :52
    loss_img = F.cross_entropy(li, lab, label_smoothing=label_smoothing, reduction="none")  # :57
    loss_txt = F.cross_entropy(lt, lab, label_smoothing=label_smoothing, reduction="none")  # :63
    loss_img = loss_img.view(B1, B2).mean(dim=1)  # :72
    loss_txt = loss_txt.view(B1, B2).mean(dim=1)  # :73
    return (loss_img + loss_txt) / 2  # :77


# --------------------------------------------------------------------------------------------
# CLOOB — reference losses/cloobLoss.py
# --------------------------------------------------------------------------------------------
def hopfield(state_patterns, stored_patterns, scale_hopfield):
    """cloobLoss.py:43-47 -> [d, N]."""
    retrieved = stored_patterns.T @ torch.softmax(
        scale_hopfield * stored_patterns @ state_patterns.T, dim=0)
    return retrieved / torch.linalg.norm(retrieved, axis=0, keepdims=True)


def hopfield_retrieval(image_features, text_features, scale_hopfield):
    """cloobLoss.py:35-40."""
    p_xx = hopfield(image_features, image_features, scale_hopfield)
    p_yy = hopfield(text_features, text_features, scale_hopfield)
    p_xy = hopfield(text_features, image_features, scale_hopfield)
    p_yx = hopfield(image_features, text_features, scale_hopfield)
    return p_xx, p_yy, p_xy, p_yx


def infoloob_loss(x, y, i, inv_tau):
    """cloobLoss.py:24-32."""
    k = x @ y.T * inv_tau
    positives = -torch.mean(torch.sum(k * i, dim=1))
    large_neg = -10000.0
    arg_lse = k * torch.logical_not(i) + i * large_neg
    negatives = torch.mean(torch.logsumexp(arg_lse, dim=1))
    return positives + negatives


def cloob_loss(image_features, text_features, inv_tau, scale_hopfield):
    """cloobLoss.py:10-21."""
    p_xx, p_yy, p_xy, p_yx = hopfield_retrieval(image_features, text_features, scale_hopfield)
    identity = torch.eye(p_xx.shape[1], dtype=image_features.dtype) > 0.5
    loss_img = infoloob_loss(p_xx.T, p_xy.T, identity, inv_tau)
    loss_txt = infoloob_loss(p_yy.T, p_yx.T, identity, inv_tau)
    return (loss_img + loss_txt) / 2


# --------------------------------------------------------------------------------------------
# AlignCLIP — reference AlignCLIP/align_clip/loss.py (single-process branch)
# --------------------------------------------------------------------------------------------
def inmodality_loss(image_features, text_features, logit_scale, semantic_features=None,
                    alpha=1.0, beta=0.5, nl_semantic_supervision=False, separate_text=True,
                    separate_image=False):
    """ClipInModalityLoss.get_logits (:209-216) + forward (:221-273).

    The only departure from the source is `torch.eye(size)` staying on the input's device:
    the original `.to(logits.get_device())` raises on CPU (get_device() == -1, :233,:237).
    Returns (total, clip_loss, inModality_loss)."""
    logits_per_image = image_features @ image_features.T  # :209
    logits_per_text = text_features @ text_features.T  # :210
    logits_image_text = image_features @ text_features.T  # :212
    ls_it = logit_scale * image_features @ text_features.T  # :215
    ls_ti = logit_scale * text_features @ image_features.T  # :216
    n = logits_per_image.shape[0]
    labels = torch.arange(n)  # :225
    if nl_semantic_supervision:
        semantic_features = semantic_features / semantic_features.norm(dim=-1, keepdim=True)  # :229
        semantic_sim = 1 - semantic_features @ semantic_features.T  # :230-231
        paired = torch.mul(logits_image_text, torch.eye(n, dtype=logits_image_text.dtype))  # :237
        terms = []
        if separate_text:  # :239-242
            lt = logit_scale * (torch.mul(logits_per_text, semantic_sim) + paired)
            terms.append(F.cross_entropy(lt, labels))
        if separate_image:  # :244-247
            li = logit_scale * (torch.mul(logits_per_image, semantic_sim) + paired)
            terms.append(F.cross_entropy(li, labels))
        in_modality = beta * sum(terms)  # :249-255
    else:  # :258-265
        in_modality = beta * (F.cross_entropy(logit_scale * logits_per_image, labels) +
                              F.cross_entropy(logit_scale * logits_per_text, labels))
    clip = alpha * (F.cross_entropy(ls_it, labels) + F.cross_entropy(ls_ti, labels))  # :267-269
    return in_modality + clip, clip, in_modality  # :272


def openclip_loss(image_features, text_features, logit_scale):
    """AlignCLIP flavour ClipLoss.forward, world_size 1 (loss.py:103-132)."""
    li = logit_scale * image_features @ text_features.T
    lt = logit_scale * text_features @ image_features.T
    labels = torch.arange(li.shape[0])
    return (F.cross_entropy(li, labels) + F.cross_entropy(lt, labels)) / 2


# --------------------------------------------------------------------------------------------
# helpers shared by tests and the bench
# --------------------------------------------------------------------------------------------
def synth_embeddings(n, d, seed, correlated=False, dtype=torch.float32):
    """seeded L2-normalised pairs (img then txt from one generator)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = F.normalize(torch.randn(n, d, generator=g, dtype=torch.float32), dim=-1)
    if correlated:
        y = F.normalize(x + 0.5 * torch.randn(n, d, generator=g, dtype=torch.float32), dim=-1)
    else:
        y = F.normalize(torch.randn(n, d, generator=g, dtype=torch.float32), dim=-1)
    return x.to(dtype), y.to(dtype)


def loss_and_grads(fn, tensors, *args, **kw):
    """Run fn(*tensors, *args) with autograd; return (loss, [grad per tensor])."""
    leaves = [t.detach().clone().requires_grad_(True) for t in tensors]
    out = fn(*leaves, *args, **kw)
    loss = out[0] if isinstance(out, tuple) else out
    loss.sum().backward()
    return loss.detach(), [l.grad for l in leaves]

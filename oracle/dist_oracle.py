"""CPU oracle of the vCLIP path — TEST INFRASTRUCTURE ONLY (imported by tests/, smoke() and the
golden generator; never by the product path).

torch-CPU restatements of the reference's distributions and of `VClipLoss`, with every random
draw passed in explicitly so the deterministic transforms can be compared exactly:
  losses/distributions/VonMisesFisher.py   (pinned: tests/golden/dist.npz made by running it)
  losses/distributions/ProjectedNormal.py  (pinned likewise)
  losses/vclipLoss.py                      (pinned likewise for vMF / ProjectedNormal / Normal)
  power_spherical @3d4619a (third party, source absent) — PARITY UNPINNED: restated from the
    published algorithm; anchored only on the reference call sites
    vclipLoss.py:12,16-20,185-189 and on self-checks in tests/test_oracle_dist.py.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from . import losses_oracle as LO


# --------------------------------------------------------------------------------------------
# von Mises-Fisher — reference losses/distributions/VonMisesFisher.py
# --------------------------------------------------------------------------------------------
def log_bessel_mpmath(nu, kappa):
    """log I_nu(kappa) the way the reference gets it (:35-37): mpmath, elementwise, float64."""
    import mpmath
    k = kappa.detach().cpu().double().reshape(-1).tolist()
    out = [float(mpmath.log(mpmath.besseli(nu, mpmath.mpf(1.0) * x))) for x in k]
    return torch.tensor(out, dtype=torch.float64).reshape(kappa.shape)


class _LogPartition(torch.autograd.Function):
    """vMFLogPartition (:7-62)."""

    @staticmethod
    def forward(ctx, d, kappa):
        s = 0.5 * d - 1
        logI = log_bessel_mpmath(s, kappa).to(kappa)  # :37
        logC = d * (-0.5 * np.log(2 * np.pi)) + s * kappa.log() - logI  # :42
        ctx.s, ctx.kappa, ctx.logI = s, kappa.detach(), logI
        return logC

    @staticmethod
    def backward(ctx, g):
        logI2 = log_bessel_mpmath(ctx.s + 1, ctx.kappa).to(ctx.logI)  # :54-55
        return None, g * (-(logI2 - ctx.logI).exp())  # :60-62


def vmf_log_partition(d, kappa):
    return _LogPartition.apply(d, kappa)


def vmf_normalize_loc(loc):
    return loc / loc.norm(dim=-1, keepdim=True).clamp(min=1e-10)  # :95


def vmf_sample(loc_unit, kappa, v, eps, u):
    """_sample_impl (:148-241) with its draws passed in: v [B,N,d-1] normals, eps [B,P] Beta
    proposals, u [B,P] uniforms (P = 10 N, the first rejection round).  Returns [N,B,d]."""
    B, d = loc_unit.shape
    N = v.shape[1]
    out = []
    for b in range(B):
        mu_b = loc_unit[b]
        kappa_b = float(kappa.reshape(-1)[b])  # .item(): no gradient to kappa (:182)
        vb = v[b] / v[b].norm(dim=1, keepdim=True).clamp(min=1e-10)  # :186
        kmr = np.sqrt(4 * kappa_b ** 2 + (d - 1) ** 2)  # :189-192
        bb = (kmr - 2 * kappa_b) / (d - 1)
        aa = (kmr + 2 * kappa_b + d - 1) / 4
        dd = (4 * aa * bb) / (1 + bb) - (d - 1) * np.log(d - 1)
        w0 = (1 - (1 + bb) * eps[b]) / (1 - (1 - bb) * eps[b])  # :212
        t0 = (2 * aa * bb) / (1 - (1 - bb) * eps[b])  # :213
        det = (d - 1) * t0.log() - t0 + dd - u[b].log()  # :214
        v0 = w0[det >= 0][:N]  # :215-218
        assert v0.numel() == N, "first rejection round did not yield N samples"
        v0 = v0.reshape(N, 1)
        sb = torch.cat([v0, (1 - v0 ** 2).sqrt() * vb], dim=1)  # :224
        e1 = torch.zeros(d, 1, dtype=mu_b.dtype)  # :227-231
        e1[0, 0] = 1.0
        e1mu = e1 - mu_b.unsqueeze(1)
        e1mu = e1mu / e1mu.norm(dim=0).clamp(min=1e-10)
        out.append(sb - 2 * (sb @ e1mu) @ e1mu.t())
    return torch.stack(out, dim=0).permute(1, 0, 2)  # :235-239


def vmf_entropy(d, kappa):
    """entropy (:299-317) with the reference's crude A_d (quirk D.6)."""
    logC = vmf_log_partition(d, kappa)
    A_d = torch.tanh(kappa) * (1 - (d - 1) / (2 * kappa + 1e-10))
    return -logC - kappa * A_d


def vmf_log_prob(loc_unit, kappa, value):
    value_norm = value / value.norm(dim=-1, keepdim=True).clamp(min=1e-10)  # :253
    dotp = (loc_unit * value_norm).sum(dim=-1)
    return vmf_log_partition(loc_unit.shape[-1], kappa) + kappa * dotp  # :259-262


# --------------------------------------------------------------------------------------------
# ProjectedNormal — reference losses/distributions/ProjectedNormal.py
# --------------------------------------------------------------------------------------------
def pn_rsample(mu, sigma, eps):
    samples = mu + eps * sigma  # Normal.rsample (:48)
    return samples / samples.norm(dim=-1, keepdim=True)  # :49


def pn_entropy(mu, sigma):
    dim = mu.size(-1)
    return 0.5 * dim * (1 + math.log(2 * math.pi)) + torch.log(sigma).sum(dim=-1)  # :89-97


# --------------------------------------------------------------------------------------------
# power_spherical (third party; restated from SURVEY App. B — parity unpinned)
# --------------------------------------------------------------------------------------------
PS_EPS = 1e-7


def ps_alpha_beta(d, scale):
    return (d - 1) / 2 + scale, torch.full_like(scale, (d - 1) / 2)


def ps_rsample(loc, scale, z, v):
    """z [S,B]: Beta(alpha, beta) draws *with the autograd graph of Beta.rsample if wanted*;
    v [S,B,d-1] normals.  Returns [S,B,d]."""
    t = (2 * z - 1).unsqueeze(-1)
    vh = v / (v.norm(dim=-1, keepdim=True) + PS_EPS)
    x = torch.cat((t, vh * torch.sqrt(torch.clamp(1 - t ** 2, PS_EPS))), -1)
    e1 = torch.zeros_like(loc)
    e1[..., 0] = 1
    u = e1 - loc
    u = u / (u.norm(dim=-1, keepdim=True) + PS_EPS)
    return x - 2 * (x * u).sum(-1, keepdim=True) * u


def ps_log_normalizer(d, scale):
    alpha, beta = ps_alpha_beta(d, scale)
    return -((alpha + beta) * math.log(2) + torch.lgamma(alpha) - torch.lgamma(alpha + beta) +
             beta * math.log(math.pi))


def ps_entropy(d, scale):
    alpha, beta = ps_alpha_beta(d, scale)
    return -(ps_log_normalizer(d, scale) +
             scale * (math.log(2) + torch.digamma(alpha) - torch.digamma(alpha + beta)))


def hu_entropy(d):
    return -(math.lgamma(d / 2) - (math.log(2) + (d / 2) * math.log(math.pi)))


def ps_kl_uniform(d, scale):
    return -ps_entropy(d, scale) + hu_entropy(d)


# --------------------------------------------------------------------------------------------
# VClipLoss — reference losses/vclipLoss.py:78-212, sampling branch, samples passed in
# --------------------------------------------------------------------------------------------
def vclip_total(image_samples, text_samples, kl_image, kl_text, image_vars, text_vars,
                logits_scale, clip_weight=1.0, kl_weight=1.0, var_reg_weight=0.1,
                label_smoothing=0.1, normal=False):
    clip = LO.clip_loss_sampled(image_samples, text_samples, logits_scale, label_smoothing).mean()
    li, lt = torch.log(image_vars + 1e-8), torch.log(text_vars + 1e-8)  # :145-146
    if normal:
        var_reg = torch.mean(((li - lt) ** 2).sum(dim=-1))  # :149-150
    else:
        var_reg = torch.mean((li - lt) ** 2)  # :152
    total = clip_weight * clip + 0.5 * kl_weight * (kl_image + kl_text) + var_reg_weight * var_reg
    return {"total_loss": total, "clip_loss": clip, "image_kl_loss": kl_image,
            "var_reg": var_reg, "text_kl_loss": kl_text}


def vmf_kl(d, kappa):
    """vclipLoss.py:192-209 (quirk D.5 surface-area expression)."""
    if d == 2:
        lsa = math.log(2 * math.pi)
    else:
        lsa = (d / 2) * math.log(2 * math.pi) - torch.lgamma(torch.tensor(d / 2.0)).item()
    return (-vmf_entropy(d, kappa) - (-lsa)).mean()

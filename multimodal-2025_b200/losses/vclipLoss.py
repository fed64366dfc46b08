"""Drop-in for the reference `losses/vclipLoss.py` (VClipLoss :43-212 and the module-level mean
helpers :11-40).  Same constructor, `forward(image_distribution, text_distribution, image_vars,
text_vars, logits_scale, is_train, kl_weight_override=None)` and the same 5-key dict.

Hot work: `rsample((S,))` of both distributions (sampler kernels, distributions/_fn.py) and the
sample-batched CLIP loss `ClipLoss.forward_sampled` (fused similarity kernels with the batch
dimension = S).  The KL / variance-regulariser terms are O(B) elementwise reductions.
"""
from __future__ import annotations

import math

import torch
from torch import nn
from torch.distributions import Distribution

from .clipLoss import ClipLoss
from .distributions.VonMisesFisher import VonMisesFisher
from .distributions.power_spherical import HypersphericalUniform, PowerSpherical


def power_spherical_mean(distribution: PowerSpherical):
    return distribution.loc * distribution.base_dist.marginal_t.mean.unsqueeze(-1)


def power_spherical_mean2(distribution: PowerSpherical):
    # the reference hard-codes d = 512 here (quirk D.9, vclipLoss.py:16-20)
    alpha = (512 - 1) / 2 + distribution.scale
    beta = (512 - 1) / 2
    mean = (alpha - beta) / (alpha + beta)
    return distribution.loc * mean.unsqueeze(-1)


def power_spherical_mean_normalized(distribution: PowerSpherical):
    return distribution.loc


def vmf_mean_normalized(distribution: VonMisesFisher):
    """Get the normalized mean (mode) of a VonMisesFisher distribution."""
    return distribution.mode


def get_distribution_mean(distribution):
    """Get the mean/mode of a distribution regardless of type (reference :32-40)."""
    if isinstance(distribution, PowerSpherical):
        return power_spherical_mean_normalized(distribution)
    elif isinstance(distribution, VonMisesFisher):
        return vmf_mean_normalized(distribution)
    else:
        return distribution.mean


class VClipLoss(nn.Module):
    """Variational CLIP loss over spherical distributions (reference :43-162)."""

    def __init__(self, clip_weight=1.0, kl_weight=1.0, num_samples=20, var_reg_weight=0.1,
                 distribution_type='power_spherical', use_mean_only=False, expected_value=False,
                 label_smoothing=0.1):
        super(VClipLoss, self).__init__()
        self.clip_weight = clip_weight
        self.kl_weight = kl_weight
        self.num_samples = num_samples
        self.distribution_type = distribution_type
        self.use_mean_only = use_mean_only
        self.var_reg_weight = var_reg_weight
        self.expected_value = expected_value
        self.clip_loss = ClipLoss(label_smoothing)

    def forward(self, image_distribution: Distribution, text_distribution: Distribution,
                image_vars: torch.Tensor, text_vars: torch.Tensor, logits_scale: torch.Tensor,
                is_train: bool, kl_weight_override: float | None = None):
        B, D = image_distribution.loc.shape
        kl_weight = kl_weight_override if kl_weight_override is not None else self.kl_weight
        use_sampling = kl_weight > 0

        if use_sampling and is_train:
            if self.expected_value:
                clip_loss = self.clip_loss.forward(power_spherical_mean2(image_distribution),
                                                   power_spherical_mean2(text_distribution),
                                                   logits_scale, normalize=False)
            else:
                image_samples = image_distribution.rsample((self.num_samples, ))
                text_samples = text_distribution.rsample((self.num_samples, ))
                clip_loss = self.clip_loss.forward_sampled(image_samples, text_samples,
                                                           logits_scale)
                clip_loss = clip_loss.mean()
        else:
            image_samples_flat = get_distribution_mean(image_distribution)
            text_samples_flat = get_distribution_mean(text_distribution)
            clip_loss = self.clip_loss.forward(image_samples_flat, text_samples_flat,
                                               logits_scale)

        kl_image = self._compute_kl_divergence(image_distribution)
        kl_text = self._compute_kl_divergence(text_distribution)

        log_image_vars = torch.log(image_vars + 1e-8)
        log_text_vars = torch.log(text_vars + 1e-8)
        if isinstance(image_distribution, torch.distributions.Normal):
            var_reg = torch.mean(((log_image_vars - log_text_vars)**2).sum(dim=-1))
        else:
            var_reg = torch.mean((log_image_vars - log_text_vars)**2)
        total_loss = self.clip_weight * clip_loss + 0.5 * kl_weight * (
            kl_image + kl_text) + self.var_reg_weight * var_reg

        return {
            'total_loss': total_loss,
            'clip_loss': clip_loss,
            'image_kl_loss': kl_image,
            'var_reg': var_reg,
            'text_kl_loss': kl_text
        }

    def _compute_kl_divergence(self, distribution):
        """KL to the uniform distribution on the sphere (reference :164-212)."""
        if isinstance(distribution, torch.distributions.Normal):
            prior = torch.distributions.Normal(torch.zeros_like(distribution.loc),
                                               torch.ones_like(distribution.scale))
            kl = torch.distributions.kl_divergence(distribution, prior)
            return kl.sum(dim=-1).mean()
        if isinstance(distribution, PowerSpherical):
            return torch.distributions.kl_divergence(
                distribution, HypersphericalUniform(dim=distribution.loc.shape[-1])).mean()
        elif isinstance(distribution, VonMisesFisher):
            d = distribution.dim
            entropy = distribution.entropy()
            if d == 2:
                log_surface_area = math.log(2 * math.pi)
            else:
                # (d/2) log(2 pi) - lgamma(d/2): the reference's expression (quirk D.5)
                log_surface_area = (d / 2) * math.log(2 * math.pi) - torch.lgamma(
                    torch.tensor(d / 2.0)).item()
            kl = -entropy - (-log_surface_area)
            return kl.mean()
        else:
            return -distribution.entropy().mean()

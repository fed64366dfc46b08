"""Drop-in for the reference `losses/vclipLoss.py` (VClipLoss :43-212 and the module-level mean
helpers :11-40).  Same constructor, `forward(image_distribution, text_distribution, image_vars,
text_vars, logits_scale, is_train, kl_weight_override=None)` and the same 5-key dict.

Hot work: `rsample((S,))` of both distributions (sampler kernels, distributions/_fn.py) and the
sample-batched CLIP loss `ClipLoss.forward_sampled` (fused similarity kernels with the batch
dimension = S).  The KL / variance-regulariser terms are O(B) elementwise reductions.

The loss is assembled from three private pieces (`_clip_term`, `_kl_to_uniform`,
`_variance_regulariser`); each evaluates the reference's expression for its term with the same
torch operations in the same order, so the values agree bit for bit on the same inputs
(tests/test_vclip_gpu.py against goldens produced by the reference).
"""
from __future__ import annotations

import math

import torch
from torch import nn
from torch.distributions import Distribution, Normal, kl_divergence

from .clipLoss import ClipLoss
from .distributions.VonMisesFisher import VonMisesFisher
from .distributions.power_spherical import HypersphericalUniform, PowerSpherical

_VAR_EPS = 1e-8          # added to the variances before the log (reference :145-146)
_HARDCODED_DIM = 512     # quirk D.9: the reference's closed-form mean assumes d = 512 (:17-18)


# ---------------------------------------------------------------------------------------------
# module-level helpers of the reference (:11-40)
# ---------------------------------------------------------------------------------------------
def power_spherical_mean(distribution: PowerSpherical):
    """E[x] = loc * E[t], E[t] taken from the marginal Beta of the distribution (:11-13)."""
    radial = distribution.base_dist.marginal_t.mean
    return distribution.loc * radial.unsqueeze(-1)


def power_spherical_mean2(distribution: PowerSpherical):
    """Closed form (alpha - beta) / (alpha + beta) with the reference's d = 512 (:16-20)."""
    half = (_HARDCODED_DIM - 1) / 2
    alpha, beta = half + distribution.scale, half
    return distribution.loc * ((alpha - beta) / (alpha + beta)).unsqueeze(-1)


def power_spherical_mean_normalized(distribution: PowerSpherical):
    """The unit-norm direction of the mean is the location itself (:23-24)."""
    return distribution.loc


def vmf_mean_normalized(distribution: VonMisesFisher):
    """Unit-norm mean direction of a von Mises-Fisher distribution: its mode (:27-29)."""
    return distribution.mode


def get_distribution_mean(distribution):
    """Point estimate used when the loss does not sample (:32-40)."""
    if isinstance(distribution, PowerSpherical):
        return power_spherical_mean_normalized(distribution)
    if isinstance(distribution, VonMisesFisher):
        return vmf_mean_normalized(distribution)
    return distribution.mean


def _log_sphere_area_reference(d: int) -> float:
    """`log S_d` exactly as the reference evaluates it (:201-206, quirk D.5: for d != 2 it uses
    (d/2) log(2 pi) - lgamma(d/2), not the true surface area)."""
    if d == 2:
        return math.log(2 * math.pi)
    return (d / 2) * math.log(2 * math.pi) - torch.lgamma(torch.tensor(d / 2.0)).item()


class VClipLoss(nn.Module):
    """Variational CLIP loss over distributions on the sphere (reference :43-162): a CLIP term
    on samples (or on point estimates), KL of both posteriors to the uniform prior, and a
    regulariser tying the two modalities' log-variances."""

    def __init__(self, clip_weight=1.0, kl_weight=1.0, num_samples=20, var_reg_weight=0.1,
                 distribution_type='power_spherical', use_mean_only=False, expected_value=False,
                 label_smoothing=0.1):
        super().__init__()
        self.clip_loss = ClipLoss(label_smoothing)
        self.clip_weight, self.kl_weight, self.var_reg_weight = clip_weight, kl_weight, var_reg_weight
        self.num_samples = num_samples
        self.distribution_type = distribution_type
        self.use_mean_only = use_mean_only
        self.expected_value = expected_value

    # -- the three terms ----------------------------------------------------------------------
    def _clip_term(self, img_dist, txt_dist, logits_scale, sample: bool):
        """(:103-132) training with a KL weight samples S embeddings per item — unless
        `expected_value` asks for the closed-form means, un-normalised; otherwise the
        distributions' point estimates go through the plain CLIP loss."""
        if not sample:
            return self.clip_loss.forward(get_distribution_mean(img_dist),
                                          get_distribution_mean(txt_dist), logits_scale)
        if self.expected_value:
            return self.clip_loss.forward(power_spherical_mean2(img_dist),
                                          power_spherical_mean2(txt_dist), logits_scale,
                                          normalize=False)
        shape = (self.num_samples,)
        per_slice = self.clip_loss.forward_sampled(img_dist.rsample(shape), txt_dist.rsample(shape),
                                                   logits_scale)
        return per_slice.mean()

    def _kl_to_uniform(self, dist):
        """KL(posterior || prior), batch mean (:164-212): N(0, 1) prior for Normal posteriors,
        the uniform distribution on the sphere otherwise."""
        if isinstance(dist, Normal):
            prior = Normal(torch.zeros_like(dist.loc), torch.ones_like(dist.scale))
            return kl_divergence(dist, prior).sum(dim=-1).mean()
        if isinstance(dist, PowerSpherical):
            prior = HypersphericalUniform(dim=dist.loc.shape[-1])
            return kl_divergence(dist, prior).mean()
        neg_entropy = -dist.entropy()
        if isinstance(dist, VonMisesFisher):
            # KL = -H(p) - log(1 / S_d) with the reference's S_d (kept term by term: :208)
            return (neg_entropy - (-_log_sphere_area_reference(dist.dim))).mean()
        return neg_entropy.mean()

    _compute_kl_divergence = _kl_to_uniform  # the reference's method name

    @staticmethod
    def _variance_regulariser(image_vars, text_vars, per_dimension: bool):
        """mean squared log-ratio of the two variances (:145-153); summed over the embedding
        dimension first when the posteriors are diagonal Normals."""
        gap = (torch.log(image_vars + _VAR_EPS) - torch.log(text_vars + _VAR_EPS)) ** 2
        return torch.mean(gap.sum(dim=-1)) if per_dimension else torch.mean(gap)

    # -- the module ---------------------------------------------------------------------------
    def forward(self, image_distribution: Distribution, text_distribution: Distribution,
                image_vars: torch.Tensor, text_vars: torch.Tensor, logits_scale: torch.Tensor,
                is_train: bool, kl_weight_override: float | None = None):
        B, D = image_distribution.loc.shape  # noqa: F841 - same shape contract as the reference
        kl_weight = self.kl_weight if kl_weight_override is None else kl_weight_override
        clip = self._clip_term(image_distribution, text_distribution, logits_scale,
                               sample=bool(kl_weight > 0 and is_train))
        kl_i = self._kl_to_uniform(image_distribution)
        kl_t = self._kl_to_uniform(text_distribution)
        reg = self._variance_regulariser(image_vars, text_vars,
                                         isinstance(image_distribution, Normal))
        total = self.clip_weight * clip + 0.5 * kl_weight * (kl_i + kl_t) + self.var_reg_weight * reg
        return {'total_loss': total, 'clip_loss': clip, 'image_kl_loss': kl_i, 'var_reg': reg,
                'text_kl_loss': kl_t}

"""B200-native stand-in for the third-party `power_spherical` package the reference imports
(`from power_spherical import PowerSpherical, HypersphericalUniform`, losses/vclipLoss.py:4;
pinned git 3d4619a9, uv.lock:1726-1728).  Its source is NOT in /root/reference, so this follows
the published algorithm as restated in SURVEY.md App. B — **parity unpinned** (DESIGN.md §2);
the constants marked there as unverified are `_EPS`.

  PowerSpherical(loc, scale):  t = 2 Beta((d-1)/2 + scale, (d-1)/2) - 1,  v ~ U(S^{d-2}),
                               x = [t, sqrt(1 - t^2) v],  y = Householder(e1 -> loc) x
Sampling runs in `mm25_beta_sample` + `mm25_sphere_sample_fwd`; the gradient w.r.t. `scale`
goes through the Beta draw by implicit reparameterisation (`mm25_beta_grad`), as
torch.distributions.Beta.rsample does for the upstream package.
"""
from __future__ import annotations

import math

import numpy as np
import torch
from torch.distributions import Distribution
from torch.distributions.kl import register_kl

from ... import ops
from ._fn import BetaReparamFn, SphereSampleFn, fresh_seed

_EPS = 1e-7


class HypersphericalUniform(Distribution):
    arg_constraints = {}
    has_rsample = False

    def __init__(self, dim, device="cpu", dtype=torch.float32, validate_args=None):
        self.dim = dim
        self.device, self.dtype = device, dtype
        super().__init__(event_shape=torch.Size([dim]), validate_args=validate_args)

    def _log_normalizer(self):
        return torch.tensor(math.lgamma(self.dim / 2) -
                            (math.log(2) + (self.dim / 2) * math.log(math.pi)),
                            device=self.device, dtype=self.dtype)

    def log_prob(self, x):
        return torch.full_like(x[..., 0], float(self._log_normalizer()))

    def entropy(self):
        return -self._log_normalizer()

    def sample(self, sample_shape=torch.Size()):
        v = torch.empty(torch.Size(sample_shape) + torch.Size([self.dim]), device=self.device,
                        dtype=self.dtype).normal_()
        return v / (v.norm(dim=-1, keepdim=True) + _EPS)


class MarginalTDistribution:
    """t = 2 Beta(alpha, beta) - 1; only what the reference touches (`.mean`, `.entropy()`,
    `.base_dist.concentration1/0`; vclipLoss.py:12)."""

    def __init__(self, dim, scale):
        self.dim, self.scale = dim, scale
        self.base_dist = torch.distributions.Beta((dim - 1) / 2 + scale,
                                                  torch.full_like(scale, (dim - 1) / 2),
                                                  validate_args=False)

    @property
    def mean(self):
        return 2 * self.base_dist.mean - 1

    def entropy(self):
        return self.base_dist.entropy() + math.log(2)


class _JointTSphericalUniform:
    def __init__(self, dim, scale):
        self.dim, self.scale = dim, scale
        self.marginal_t = MarginalTDistribution(dim, scale)


class PowerSpherical(Distribution):
    arg_constraints = {}
    has_rsample = True

    def __init__(self, loc, scale, validate_args=None):
        self.loc, self.scale = loc, scale
        self.dim = loc.shape[-1]
        self.base_dist = _JointTSphericalUniform(self.dim, scale)
        self._injected_noise = None  # parity hook: {"z": [S,B] Beta draws, "v": [S,B,d-1] normals}
        super().__init__(batch_shape=loc.shape[:-1], event_shape=torch.Size([self.dim]),
                         validate_args=validate_args)

    def _alpha_beta(self):
        beta = torch.full_like(self.scale, (self.dim - 1) / 2)
        return (self.dim - 1) / 2 + self.scale, beta

    def rsample(self, sample_shape=torch.Size()):
        sample_shape = torch.Size(sample_shape)
        S = int(np.prod(sample_shape)) if len(sample_shape) > 0 else 1
        d = self.dim
        loc2 = self.loc.reshape(-1, d)
        alpha, beta = self._alpha_beta()
        a1, b1 = alpha.reshape(-1), beta.reshape(-1)
        seed = fresh_seed()
        noise = self._injected_noise
        if noise is not None:
            z, v_in = noise["z"].reshape(S, -1).float(), noise["v"]
        else:
            z, v_in = ops.beta_sample(a1.detach(), b1.detach(), S, seed), None
        z = BetaReparamFn.apply(z, a1, b1)
        t = 2 * z - 1
        y = SphereSampleFn.apply(t, loc2, v_in, seed, ops.POWER_SPHERICAL_CFG)
        return y.reshape(sample_shape + self.loc.shape)

    def sample(self, sample_shape=torch.Size()):
        with torch.no_grad():
            return self.rsample(sample_shape)

    def log_normalizer(self):
        alpha, beta = self._alpha_beta()
        return -((alpha + beta) * math.log(2) + torch.lgamma(alpha) - torch.lgamma(alpha + beta) +
                 beta * math.log(math.pi))

    def log_prob(self, value):
        return self.log_normalizer() + self.scale * torch.log1p((self.loc * value).sum(-1))

    def entropy(self):
        alpha, beta = self._alpha_beta()
        return -(self.log_normalizer() + self.scale *
                 (math.log(2) + torch.digamma(alpha) - torch.digamma(alpha + beta)))

    @property
    def mean(self):
        return self.loc * self.base_dist.marginal_t.mean.unsqueeze(-1)


@register_kl(PowerSpherical, HypersphericalUniform)
def _kl_powerspherical_uniform(p, q):
    return -p.entropy() + q.entropy().to(p.loc.device)

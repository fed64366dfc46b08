"""Drop-in for the reference `losses/distributions/ProjectedNormal.py` (lines 1-118):
`ProjectedNormal(mu, sigma, validate_args=None)` with `sample / rsample / log_prob / entropy /
mean / mode / loc`.  `rsample` is one warp-per-row kernel (`mm25_sample_projnormal_fwd`) with a
fused backward that reduces the gradients of `mu` and `sigma` over the samples."""
from __future__ import annotations

import math

import numpy as np
import torch
from torch.distributions import Distribution

from ._fn import ProjNormalSampleFn, RowNormFn, fresh_seed


class ProjectedNormal(Distribution):
    arg_constraints = {}

    def __init__(self, mu, sigma, validate_args=None):
        self.mu = mu
        self.sigma = sigma
        # normalised mean direction, no eps (reference :25)
        self._mean_direction = RowNormFn.apply(mu, 0.0)
        super().__init__(validate_args=validate_args)
        self.underlying_normal = torch.distributions.Normal(mu, sigma)
        self.has_rsample = True
        self._injected_noise = None  # parity hook: {"eps": [S, B, d] standard normals}

    def _draw(self, sample_shape):
        sample_shape = torch.Size(sample_shape)
        S = int(np.prod(sample_shape)) if len(sample_shape) > 0 else 1
        d = self.mu.shape[-1]
        mu2 = self.mu.reshape(-1, d)
        sg2 = self.sigma.expand_as(self.mu).reshape(-1, d)
        eps = None if self._injected_noise is None else self._injected_noise["eps"]
        out = ProjNormalSampleFn.apply(mu2, sg2, eps, fresh_seed(), S)
        return out.reshape(sample_shape + self.mu.shape)

    def sample(self, sample_shape=torch.Size()):
        with torch.no_grad():
            return self._draw(sample_shape)

    def rsample(self, sample_shape=torch.Size()):
        return self._draw(sample_shape)

    def log_prob(self, value):
        """reference :51-87 (its simplified density)."""
        value_norm = value / value.norm(dim=-1, keepdim=True)
        mu_norm = self.mu.norm(dim=-1, keepdim=True)
        dot_product = (self.mu * value_norm).sum(dim=-1, keepdim=True)
        log_sigma_sq = 2 * torch.log(self.sigma)
        exponent = -0.5 * (mu_norm ** 2 / (self.sigma ** 2)).sum(dim=-1)
        exponent = exponent + 0.5 * ((dot_product ** 2) / (self.sigma ** 2)).sum(dim=-1)
        dim = self.mu.size(-1)
        log_norm = -0.5 * dim * math.log(2 * math.pi) - 0.5 * log_sigma_sq.sum(dim=-1)
        return log_norm + exponent

    def entropy(self):
        """Gaussian entropy 0.5 d (1 + log 2 pi) + sum log sigma (reference :89-97)."""
        dim = self.mu.size(-1)
        log_sigma_sum = torch.log(self.sigma).sum(dim=-1)
        return 0.5 * dim * (1 + math.log(2 * math.pi)) + log_sigma_sum

    @property
    def mean(self):
        return self._mean_direction

    @property
    def mode(self):
        return self._mean_direction

    @property
    def loc(self):
        return self._mean_direction

"""Drop-in for the reference `losses/distributions/VonMisesFisher.py`.

Same class / method surface (`VonMisesFisher(loc, concentration, validate_args=None)` with
`sample / rsample / log_prob / entropy / mean / mode / variance / expand / dim`, and
`vMFLogPartition`).  What changes is where the work runs:

* `_sample_impl` (reference :148-241, a Python loop over batch rows with `.item()` syncs) is one
  scalar kernel for Wood's rejection sampler (`mm25_vmf_sample_w`) plus one warp-per-row kernel
  for `[w, sqrt(1 - w^2) v]` and the Householder reflection (`mm25_sphere_sample_fwd`).  As in
  the reference the sample is not reparameterised in kappa (quirk D.7): gradient reaches `loc`
  only, through the reflection.
* `vMFLogPartition` (reference :7-62, mpmath on the CPU) evaluates log I_nu(kappa) on the device
  (`mm25_log_bessel`, fp64 uniform asymptotic expansion).
"""
from __future__ import annotations

import numpy as np
import torch
from torch.distributions import Distribution

from ... import ops
from ._fn import RowNormFn, SphereSampleFn, VMFLogPartitionFn, fresh_seed


class vMFLogPartition(VMFLogPartitionFn):
    """log C_d(kappa); `vMFLogPartition.apply(d, kappa)` as in the reference (:7-62)."""


class VonMisesFisher(Distribution):
    """von Mises-Fisher distribution on the unit sphere (reference :65-316)."""

    arg_constraints = {}

    def __init__(self, loc, concentration, validate_args=None):
        # Normalize loc to unit length (reference :95)
        self.loc = RowNormFn.apply(loc, 1e-10)
        self.concentration = concentration
        self.dim = self.loc.shape[-1]
        if self.concentration.dim() < self.loc.dim() - 1:  # reference :102-105
            self.concentration = self.concentration.view(
                self.concentration.shape + (1,) * (self.loc.dim() - self.concentration.dim() - 1))
        batch_shape = self.loc.shape[:-1]
        event_shape = torch.Size([self.dim])
        # parity hook: {"v": [S,B,d-1] normals, "eps": [B,P] Beta proposals, "u": [B,P] uniforms}
        self._injected_noise = None
        super().__init__(batch_shape, event_shape, validate_args=validate_args)

    def expand(self, batch_shape, _instance=None):
        new = self._get_checked_instance(VonMisesFisher, _instance)
        batch_shape = torch.Size(batch_shape)
        new.loc = self.loc.expand(batch_shape + (self.dim,))
        new.concentration = self.concentration.expand(batch_shape)
        new.dim = self.dim
        new._injected_noise = None
        super(VonMisesFisher, new).__init__(batch_shape, self.event_shape, validate_args=False)
        new._validate_args = self._validate_args
        return new

    def sample(self, sample_shape=torch.Size()):
        with torch.no_grad():
            return self._sample_impl(sample_shape)

    def rsample(self, sample_shape=torch.Size()):
        return self._sample_impl(sample_shape)

    def _sample_impl(self, sample_shape=torch.Size()):
        sample_shape = torch.Size(sample_shape)
        shape = self._extended_shape(sample_shape)
        N = int(np.prod(sample_shape)) if len(sample_shape) > 0 else 1
        if N == 0:
            return torch.empty(shape, device=self.loc.device, dtype=self.loc.dtype)
        d = self.dim
        loc_flat = self.loc.reshape(-1, d)
        if self.concentration.dim() > 1:
            conc_flat = self.concentration.squeeze(-1).reshape(-1)
        else:
            conc_flat = self.concentration.reshape(-1)
        batch_size = loc_flat.shape[0]
        if conc_flat.shape[0] != batch_size:
            raise ValueError(f"Concentration batch size {conc_flat.shape[0]} doesn't match loc "
                             f"batch size {batch_size}")
        conc_flat = conc_flat.detach()  # kappa_b = conc_flat[b].item(): no gradient (:182)
        noise = self._injected_noise
        seed = fresh_seed()
        if noise is not None:
            w, count = ops.vmf_sample_w(conc_flat, d, N, seed, noise["eps"], noise["u"])
            v_in = noise["v"]
        else:
            w, count = ops.vmf_sample_w(conc_flat, d, N, seed, None, None)
            v_in = None
        samples = SphereSampleFn.apply(w, loc_flat, v_in, seed, ops.VMF_CFG)  # [N, B, d]
        target_shape = sample_shape + self.batch_shape + self.event_shape
        return samples.reshape(target_shape)

    def log_prob(self, value):
        if self._validate_args:
            self._validate_sample(value)
        value_norm = value / value.norm(dim=-1, keepdim=True).clamp(min=1e-10)
        dotp = (self.loc * value_norm).sum(dim=-1)
        logC = vMFLogPartition.apply(self.dim, self.concentration)
        return logC + self.concentration * dotp

    @property
    def mean(self):
        print("Warning: mean() currently returns mode (loc) as an approximation.")
        return self.loc

    @property
    def mode(self):
        return self.loc

    @property
    def variance(self):
        return 1.0 / (self.concentration.unsqueeze(-1) + 1e-10)

    def entropy(self):
        """H = -log C_d(kappa) - kappa A_d(kappa) with the reference's crude A_d (:299-317)."""
        logC = vMFLogPartition.apply(self.dim, self.concentration)
        A_d = torch.tanh(self.concentration) * (
            1 - (self.dim - 1) / (2 * self.concentration + 1e-10))
        return -logC - self.concentration * A_d

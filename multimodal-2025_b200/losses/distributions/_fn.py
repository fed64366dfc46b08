"""autograd.Function wrappers of the sampler kernels (ops `mm25::projnormal_*`, `sphere_*`,
`beta_*`, `log_bessel`, `rownorm_*`).  Shared by the distribution classes of this package."""
from __future__ import annotations

import torch

from ... import ops


def fresh_seed() -> int:
    """A 63-bit Philox key drawn from torch's default CPU generator: reproducible under
    torch.manual_seed, no device synchronisation.  In a multi-process run every rank usually
    calls torch.manual_seed with the SAME seed (AlignCLIP/main/train.py does), and the sampler
    kernels index their counters by LOCAL row — so the rank is folded into the key
    (splitmix64 of the draw and the rank): the rows of a sharded global batch then see
    independent noise instead of P copies of the same stream."""
    key = int(torch.empty((), dtype=torch.int64).random_().item())
    try:
        import torch.distributed as dist
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    except Exception:  # noqa: BLE001 - no process group: single-process key
        rank = 0
    if rank:
        z = (key + 0x9E3779B97F4A7C15 * (rank + 1)) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        key = (z ^ (z >> 31)) & 0x7FFFFFFFFFFFFFFF
    return key


class RowNormFn(torch.autograd.Function):
    """x / max(||x||, eps) over the last dim of a [B, d] matrix (kernels of prep.cu)."""

    @staticmethod
    def forward(ctx, x, eps):
        shp = x.shape
        x2 = x.detach().float().reshape(1, -1, shp[-1])
        xhat, inv, _ = ops.rownorm_fwd(x2, True, float(eps), ops.F32, False)
        ctx.save_for_backward(xhat, inv)
        ctx.shp, ctx.dt = shp, x.dtype
        return xhat.reshape(shp).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        xhat, inv = ctx.saved_tensors
        g3 = g.detach().float().reshape(xhat.shape).contiguous()
        dx = ops.rownorm_bwd(g3, xhat, inv, True, ops.F32, None)
        return dx.reshape(ctx.shp).to(ctx.dt), None


class ProjNormalSampleFn(torch.autograd.Function):
    """samples [S, B, d] = normalize(mu + sigma * eps)  (ProjectedNormal.py:41-49)."""

    @staticmethod
    def forward(ctx, mu, sigma, eps_in, seed, S):
        out, inv = ops.projnormal_fwd(mu, sigma, eps_in, seed, S)
        ctx.save_for_backward(out, inv, eps_in if eps_in is not None else torch.empty(0))
        ctx.seed, ctx.has_eps = seed, eps_in is not None
        ctx.dts = (mu.dtype, sigma.dtype)
        return out.to(mu.dtype)

    @staticmethod
    def backward(ctx, g):
        out, inv, eps = ctx.saved_tensors
        dmu, dsg = ops.projnormal_bwd(g.contiguous(), out, inv, eps if ctx.has_eps else None,
                                      ctx.seed)
        return dmu.to(ctx.dts[0]), dsg.to(ctx.dts[1]), None, None, None


class SphereSampleFn(torch.autograd.Function):
    """y [S, B, d] = Householder(e1 -> loc) [t, sqrt(1 - t^2) v / |v|]; grads to t and loc."""

    @staticmethod
    def forward(ctx, t, loc, v_in, seed, cfg):
        y = ops.sphere_fwd(t, loc, v_in, seed, *cfg)
        ctx.save_for_backward(y, t.detach().float(), loc.detach().float())
        ctx.cfg, ctx.dts = cfg, (t.dtype, loc.dtype)
        return y.to(loc.dtype)

    @staticmethod
    def backward(ctx, g):
        y, t, loc = ctx.saved_tensors
        dloc, dt = ops.sphere_bwd(g.contiguous(), y, t, loc, *ctx.cfg)
        return (dt.to(ctx.dts[0]) if ctx.needs_input_grad[0] else None,
                dloc.to(ctx.dts[1]) if ctx.needs_input_grad[1] else None, None, None, None)


class BetaReparamFn(torch.autograd.Function):
    """Makes Beta(alpha, beta) draws z [S, B] differentiable w.r.t. alpha [B] by implicit
    reparameterisation (what torch.distributions.Beta.rsample's backward computes)."""

    @staticmethod
    def forward(ctx, z, alpha, beta):
        ctx.save_for_backward(z, alpha.detach(), beta.detach())
        ctx.adt = alpha.dtype
        return z.clone()

    @staticmethod
    def backward(ctx, g):
        z, alpha, beta = ctx.saved_tensors
        dz = ops.beta_grad(z, alpha, beta)
        return None, (g.float() * dz).sum(dim=0).reshape(alpha.shape).to(ctx.adt), None


class VMFLogPartitionFn(torch.autograd.Function):
    """log C_d(kappa) of the vMF density with d/dkappa = -I_{nu+1}/I_nu, nu = d/2 - 1
    (reference vMFLogPartition, VonMisesFisher.py:7-62).  The two log-Bessel values are computed
    in fp64 on the device and rounded to kappa's dtype exactly where the reference casts its
    mpmath results (:37,55), so the derivative reproduces the reference's exp(logI2 - logI)."""

    nhlog2pi = -0.9189385332046727  # -0.5 * log(2 pi)

    @staticmethod
    def forward(ctx, d, kappa):
        s = 0.5 * d - 1
        logI64, logI2_64 = ops.log_bessel(kappa, float(s))
        logI = logI64.to(kappa.dtype).reshape(kappa.shape)
        logC = d * VMFLogPartitionFn.nhlog2pi + s * kappa.log() - logI
        ctx.save_for_backward(logI, logI2_64.to(kappa.dtype).reshape(kappa.shape))
        return logC

    @staticmethod
    def backward(ctx, grad_output):
        logI, logI2 = ctx.saved_tensors
        return None, grad_output * (-(logI2 - logI).exp())

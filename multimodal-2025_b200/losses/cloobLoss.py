"""Drop-in for the reference `losses/cloobLoss.py` (CLOOB: Hopfield retrieval + InfoLOOB,
https://arxiv.org/pdf/2110.11316).  Same free functions and `CLOOBLoss(inv_tau, scale_hopfield,
device)` module; nothing N x N is materialised:

  hopfield(Q, K, beta)   = rownorm( softmax_rows(beta Q K^T) K )^T      (cloobLoss.py:43-47)
  infoloob_loss(x, y, .) = -mean_i k_ii + mean_i LSE_j k'_ij,  k = inv_tau x y^T,
                           k'_ii = -10000                                (cloobLoss.py:24-32)
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops

_OPS = ops  # default backend of the autograd functions below (they accept a stand-in)


def _scalar(v, dev):
    # a fill kernel, not a host->device copy: stays legal under CUDA-graph capture
    return torch.full((1,), float(v), dtype=torch.float32, device=dev)


def _tc_operands(x, y):
    """bf16 copies for the tensor-core kernels when the inputs are bf16 (or COMPUTE_DTYPE asks
    for it) and the shape is eligible; else None (fp32 SIMT kernels, the reference's dtype)."""
    want = (x.dtype == torch.bfloat16 and y.dtype == torch.bfloat16) or \
        COMPUTE_DTYPE == torch.bfloat16
    if not (want and x.is_cuda):
        return None
    xb = x.detach().to(torch.bfloat16).contiguous()
    yb = y.detach().to(torch.bfloat16).contiguous()
    return (xb, yb) if ops.hopfield_tc_eligible(xb, yb) and ops.hopfield_tc_eligible(yb, xb) \
        else None


# None: compute in fp32 on the SIMT kernels unless the inputs are bf16 (reference semantics, rel
# 1e-5 / 1e-4).  torch.bfloat16: fp32 inputs are rounded to bf16 once per GEMM operand and run the
# tcgen05 kernels (bf16 contract 2e-3 / 2e-2), as under torch.autocast(dtype=bfloat16).
COMPUTE_DTYPE = None


def _split_ok(ops, x, y, hopfield):
    """fp32 inputs keep the reference's fp32-grade arithmetic on the tensor cores when the shape
    is covered (two-term bf16 splits, include/mm25.h mm25_split_bf16x2); MM25_FP32_TC=0 or a
    stand-in backend (CPU tests) take the fp32 FFMA / dense route."""
    import os
    if not (x.is_cuda and hasattr(ops, "split_bf16x2")) or os.environ.get("MM25_FP32_TC", "1") == "0":
        return False
    if x.dtype == torch.bfloat16 or COMPUTE_DTYPE == torch.bfloat16 or x.dim() != 2:
        return False
    M, d = x.shape
    N = y.shape[0]
    return bool(ops.hopfield_split_eligible(M, N, d) if hopfield
                else ops.sim_split_eligible(1, M, N, d))


class _HopfieldFn(torch.autograd.Function):
    """U[Nq, d] = rownorm(softmax(beta Q K^T) K)."""

    @staticmethod
    def forward(ctx, q, k, beta, backend=None):
        # `backend`: the ops module (default) or a stand-in with the same functions — the gloo
        # tests of dist.ShardedCLOOBLoss exercise this host logic on CPU through it
        ops = backend if backend is not None else _OPS
        ctx.ops = ops
        in_dtype = q.dtype
        tc = _tc_operands(q, k)
        if tc is not None:
            # tensor cores: row LSE from the PLAIN sweep; O = P K from the gradient-panel kernel
            # with w_row = 1 (bf16 P panel) and the grouped GEMM
            qb, kb = tc
            sc = _scalar(beta, q.device)
            q3, k3 = qb.unsqueeze(0), kb.unsqueeze(0)
            lse, _, _, _ = ops.sim_fwd(q3, k3, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                       ops.SIM_PLAIN, 0, False, ops.IMPL_TC)
            o, _, _, _ = ops.sim_bwd(q3, k3, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                     ops.SIM_PLAIN, 0, lse, None, 1.0, 0.0, 0.0, 0.0, None, True,
                                     False, False, ops.IMPL_TC)
            o = o / float(beta)
            u, inv, _ = ops.rownorm_fwd(o, True, 1e-30, ops.F32, False)
            ctx.save_for_backward(qb, kb, o[0], u[0], inv, lse[0], sc)
            ctx.beta, ctx.in_dtype, ctx.tc = float(beta), in_dtype, True
            return u[0].to(in_dtype)
        ctx.tc = False
        qf = q.detach().float().contiguous()
        kf = k.detach().float().contiguous()
        sc = _scalar(beta, q.device)
        if _split_ok(ops, qf, kf, True):
            # fp32-grade on the tensor cores: split operands swept over K = 3 d
            q4, qh, ql = ops.split_bf16x2(qf.unsqueeze(0), 0)
            k4, kh, kl = ops.split_bf16x2(kf.unsqueeze(0), 1)
            lse, _, _, _ = ops.sim_fwd(q4, k4, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                       ops.SIM_PLAIN, 0, False, ops.IMPL_AUTO)
            o, _, _ = ops.sim_bwd_split(q4, k4, qh, ql, kh, kl, sc, ops.SCALE_MUL, 0.0, 0, lse,
                                        None, 1.0, 0.0, 0.0, 0.0, None, True, False, 0.0)
            o = o / float(beta)
            u, inv, _ = ops.rownorm_fwd(o, True, 1e-30, ops.F32, False)
            ctx.save_for_backward(q4[0], k4[0], qh[0], ql[0], kh[0], kl[0], o[0], u[0], inv,
                                  lse[0], sc)
            ctx.beta, ctx.in_dtype, ctx.tc = float(beta), in_dtype, "split"
            return u[0].to(in_dtype)
        q3, k3 = qf.unsqueeze(0), kf.unsqueeze(0)
        lse, _, _, _ = ops.sim_fwd(q3, k3, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                   ops.SIM_PLAIN, 0, False, ops.IMPL_SIMT)
        # softmax-weighted sum = row pass with G = exp(z - lse); the kernel returns beta * O
        o, _, _, _ = ops.sim_bwd(q3, k3, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                 ops.SIM_PLAIN, 0, lse, None, 1.0, 0.0, 0.0, 0.0, None, True,
                                 False, False, ops.IMPL_SIMT)
        o = o / float(beta)
        u, inv, _ = ops.rownorm_fwd(o, True, 1e-30, ops.F32, False)
        ctx.save_for_backward(qf, kf, o[0], u[0], inv, lse[0])
        ctx.beta = float(beta)
        ctx.in_dtype = in_dtype
        return u[0].to(in_dtype)

    @staticmethod
    def backward(ctx, du):
        ops = ctx.ops
        if ctx.tc == "split":
            q4, k4, qh, ql, kh, kl, o, u, inv, lse, sc = ctx.saved_tensors
            du = du.detach().float().contiguous()
            do = ops.rownorm_bwd(du.unsqueeze(0), u.unsqueeze(0), inv, True, ops.F32, None)[0]
            delta = ops.rowdot(do, o)
            dq, dk = ops.hopfield_bwd_split(q4, k4, qh, ql, kh, kl, do, sc, lse, delta,
                                            ctx.needs_input_grad[0], ctx.needs_input_grad[1])
            return (dq.to(ctx.in_dtype) if ctx.needs_input_grad[0] else None,
                    dk.to(ctx.in_dtype) if ctx.needs_input_grad[1] else None, None, None)
        if ctx.tc:
            qb, kb, o, u, inv, lse, sc = ctx.saved_tensors
            du = du.detach().float().contiguous()
            do = ops.rownorm_bwd(du.unsqueeze(0), u.unsqueeze(0), inv, True, ops.F32, None)[0]
            delta = ops.rowdot(do, o)
            dq, dk = ops.hopfield_bwd_tc(qb, kb, do, sc, lse, delta, ctx.needs_input_grad[0],
                                         ctx.needs_input_grad[1])
            return (dq.to(ctx.in_dtype) if ctx.needs_input_grad[0] else None,
                    dk.to(ctx.in_dtype) if ctx.needs_input_grad[1] else None, None, None)
        qf, kf, o, u, inv, lse = ctx.saved_tensors
        beta = ctx.beta
        du = du.detach().float().contiguous()
        do = ops.rownorm_bwd(du.unsqueeze(0), u.unsqueeze(0), inv, True, ops.F32, None)[0]
        D = ops.rowdot(do, o)
        dq = dk = None
        if ctx.needs_input_grad[0]:
            dq = ops.hopfield_bwd_pass(qf, kf, do, kf, kf, None, beta, lse, D, False)
            dq = dq.to(ctx.in_dtype)
        if ctx.needs_input_grad[1]:
            dk = ops.hopfield_bwd_pass(kf, qf, kf, do, qf, do, beta, lse, D, True)
            dk = dk.to(ctx.in_dtype)
        return dq, dk, None, None


class _InfoLoobFn(torch.autograd.Function):
    """InfoLOOB of the rows `x` [M, d] against the columns `y` [N, d]; the positive of row i is
    column i + diag_offset and the means run over n_norm rows (defaults: square problem, M).
    A row shard of a larger problem passes its offset and the global row count and gets its
    additive part of the loss (dist.ShardedCLOOBLoss)."""

    @staticmethod
    def forward(ctx, x, y, inv_tau, diag_offset=0, n_norm=None, backend=None):
        ops = backend if backend is not None else _OPS
        ctx.ops = ops
        in_dtype = x.dtype
        tc = _tc_operands(x, y)
        impl = ops.IMPL_SIMT if tc is None else ops.IMPL_TC
        ctx.split = None
        if tc is None:
            xf = x.detach().float().contiguous().unsqueeze(0)
            yf = y.detach().float().contiguous().unsqueeze(0)
            if _split_ok(ops, xf[0], yf[0], False):
                x4, xh, xl = ops.split_bf16x2(xf, 0)
                y4, yh, yl = ops.split_bf16x2(yf, 1)
                ctx.split = (xh, xl, yh, yl)
                xf, yf, impl = x4, y4, ops.IMPL_AUTO
        else:
            xf, yf = tc[0].unsqueeze(0), tc[1].unsqueeze(0)
        sc = _scalar(inv_tau, x.device)
        lse, _, _, diag = ops.sim_fwd(xf, yf, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                      ops.SIM_MASKDIAG, int(diag_offset), False, impl)
        n = xf.shape[1] if n_norm is None else int(n_norm)
        ctx.off, ctx.n_norm = int(diag_offset), n
        # -mean(diag) + mean(lse): one deterministic reduction kernel (weights 1/2n * 2)
        loss = ops.clip_loss_finalize(lse * 2.0, diag, None, None, None, n, sc, ops.SCALE_MUL,
                                      0.0, 0.0)
        ctx.save_for_backward(xf, yf, lse, sc)
        ctx.in_dtype = in_dtype
        ctx.impl = impl
        return loss.reshape(())

    @staticmethod
    def backward(ctx, gout):
        ops = ctx.ops
        xf, yf, lse, sc = ctx.saved_tensors
        n = float(ctx.n_norm)
        g = gout.detach().float().reshape(1).contiguous()
        if ctx.split is not None:
            xh, xl, yh, yl = ctx.split
            da, db, _ = ops.sim_bwd_split(xf, yf, xh, xl, yh, yl, sc, ops.SCALE_MUL, 0.0, ctx.off,
                                          lse, None, 1.0 / n, 0.0, 1.0 / n, 0.0, g,
                                          ctx.needs_input_grad[0], ctx.needs_input_grad[1], 0.0,
                                          ops.SIM_MASKDIAG)
            dx = da[0].to(ctx.in_dtype) if ctx.needs_input_grad[0] else None
            dy = db[0].to(ctx.in_dtype) if ctx.needs_input_grad[1] else None
            return dx, dy, None, None, None, None
        da, db, _, _ = ops.sim_bwd(xf, yf, None, None, None, sc, ops.SCALE_MUL, 0.0,
                                   ops.SIM_MASKDIAG, ctx.off, lse, None, 1.0 / n, 0.0, 1.0 / n,
                                   0.0, g, ctx.needs_input_grad[0], ctx.needs_input_grad[1],
                                   False, ctx.impl)
        dx = da[0].to(ctx.in_dtype) if ctx.needs_input_grad[0] else None
        dy = db[0].to(ctx.in_dtype) if ctx.needs_input_grad[1] else None
        return dx, dy, None, None, None, None


def hopfield(state_patterns, stored_patterns, scale_hopfield):
    """Retrieved, column-normalised patterns [d, N_state] (reference cloobLoss.py:43-47)."""
    return _HopfieldFn.apply(state_patterns, stored_patterns, float(scale_hopfield), None).T


def hopfield_retrieval(image_features, text_features, scale_hopfield):
    """reference cloobLoss.py:35-40."""
    patterns_xx = hopfield(state_patterns=image_features, stored_patterns=image_features,
                           scale_hopfield=scale_hopfield)
    patterns_yy = hopfield(state_patterns=text_features, stored_patterns=text_features,
                           scale_hopfield=scale_hopfield)
    patterns_xy = hopfield(state_patterns=text_features, stored_patterns=image_features,
                           scale_hopfield=scale_hopfield)
    patterns_yx = hopfield(state_patterns=image_features, stored_patterns=text_features,
                           scale_hopfield=scale_hopfield)
    return patterns_xx, patterns_yy, patterns_xy, patterns_yx


def infoloob_loss(x, y, i, inv_tau, device=None):
    """InfoLOOB (reference cloobLoss.py:24-32).  `i` is the positives mask of the reference
    signature; the fused kernel assumes it is the identity (as cloob_loss passes it) and never
    reads it, so no N x N mask has to exist (pass None)."""
    return _InfoLoobFn.apply(x, y, float(inv_tau), 0, None, None)


def cloob_loss(image_features, text_features, inv_tau, scale_hopfield, device=None):
    """reference cloobLoss.py:10-21."""
    p_xx, p_yy, p_xy, p_yx = hopfield_retrieval(image_features, text_features, scale_hopfield)
    loss_img = infoloob_loss(p_xx.T, p_xy.T, None, inv_tau=inv_tau, device=device)
    loss_txt = infoloob_loss(p_yy.T, p_yx.T, None, inv_tau=inv_tau, device=device)
    return (loss_img + loss_txt) / 2


class CLOOBLoss(nn.Module):

    def __init__(self, inv_tau, scale_hopfield, device):
        super().__init__()
        self.inv_tau = inv_tau
        self.scale_hopfield = scale_hopfield
        self.device = device

    def forward(self, image_features, text_features):
        return cloob_loss(
            image_features, text_features, self.inv_tau, self.scale_hopfield, self.device
        )

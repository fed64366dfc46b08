"""Drop-in for the reference `losses/clipLoss.py` (ClipLoss, lines 13-121).

Same constructor, `forward(image_features, text_features, logits_scale, normalize=True)`,
`forward_sampled(image_features[S,B,E], text_features[S,B,E], logits_scale)` and the `smoothing`
buffer; the math runs in fused sm_100a kernels that never materialise the B x B logits:

    L = (sum_i r_i + sum_j c_j) / 2N - (1-eps)/N sum_i z_ii - eps/N^2 sum_ij z_ij
    r_i = LSE_j z_ij, c_j = LSE_i z_ij, z = s * xhat . yhat^T          

which equals the reference's 0.5 * (CE_eps(Z, I) + CE_eps(Z^T, I)) (clipLoss.py:116-119).
"""
from __future__ import annotations

import os

import torch
from torch import nn

from .. import ops

_EPS_NORM = 1e-12  # F.normalize default eps (clipLoss.py:95)


class _ClipLossFn(torch.autograd.Function):
    """loss[B] of B independent CLIP problems; inputs [B, M, d]."""

    @staticmethod
    def forward(ctx, img, txt, logits_scale, eps_ls, normalize, clamp_max, scale_kind, impl,
                compute_bf16=False):
        in_dtype = img.dtype
        if in_dtype not in (torch.float32, torch.bfloat16):
            img, txt = img.float(), txt.float()
        if txt.dtype != img.dtype:
            txt = txt.to(img.dtype)
        # compute_bf16: fp32 inputs are normalised in fp32 and rounded to bf16 ONCE for the
        # tensor-core kernels (what autocast does to the reference's matmul); statistics, loss
        # and gradients stay fp32.  Default: compute in the input dtype (fp32 -> 1e-5 contract).
        cdt = ops.BF16 if (img.dtype == torch.bfloat16 or compute_bf16) else ops.F32
        want_cs = eps_ls != 0.0
        xhat, invx, sx = ops.rownorm_fwd(img, normalize, _EPS_NORM, cdt, want_cs)
        yhat, invy, sy = ops.rownorm_fwd(txt, normalize, _EPS_NORM, cdt, want_cs)
        scale32 = logits_scale.detach().reshape(1).to(torch.float32).contiguous()
        B, M, _ = xhat.shape
        bound = ops.DOT_BOUND_NORMALIZED if normalize else 0.0
        N = yhat.shape[1]
        # fp32 inputs keep the reference's fp32-grade arithmetic (clipLoss.py:106 `.float()`) ON
        # THE TENSOR CORES when the shape allows: two-term bf16 splits of the normalised rows
        # (16 significant bits) swept over K = 3 d (include/mm25.h mm25_split_bf16x2); otherwise
        # the fp32 FFMA kernels.  MM25_FP32_TC=0 forces the FFMA kernels.
        split = (cdt == ops.F32 and impl == ops.IMPL_AUTO and xhat.is_cuda
                 and os.environ.get("MM25_FP32_TC", "1") != "0"
                 and ops.sim_split_eligible(B, M, N, xhat.shape[-1]))
        ctx.split = None
        e_keep = e_flag = None
        if split:
            x4, xh, xl = ops.split_bf16x2(xhat, 0)
            y4, yh, yl = ops.split_bf16x2(yhat, 1)
            row_lse, cmax, csum, diag = ops.sim_fwd(x4, y4, None, None, None, scale32,
                                                    scale_kind, clamp_max, ops.SIM_PLAIN, 0,
                                                    True, impl, bound)
            ctx.split = (x4, y4, xh, xl, yh, yl)
        else:
            # Keep E = exp(z - shift) (bf16, the activation autograd would have saved) so the
            # backward forms G' from it instead of recomputing the S tiles (mm25.h e_keep).
            keep = (any(ctx.needs_input_grad[:3]) and bound > 0.0
                    and impl != ops.IMPL_SIMT and ops.tc_eligible_bwd(xhat, yhat)
                    and 2 * B * M * N <= ops.keep_e_bytes_limit(M, N, xhat.shape[-1]))
            if keep:
                row_lse, cmax, csum, diag, e_keep, e_flag = ops.sim_fwd_keep(
                    xhat, yhat, scale32, scale_kind, clamp_max, 0, True, impl, bound)
            else:
                row_lse, cmax, csum, diag = ops.sim_fwd(xhat, yhat, None, None, None, scale32,
                                                        scale_kind, clamp_max, ops.SIM_PLAIN, 0,
                                                        True, impl, bound)
        col_lse = ops.col_merge(cmax.reshape(1, -1), csum.reshape(1, -1)).reshape(B, -1)
        loss = ops.clip_loss_finalize(row_lse, diag, col_lse, sx if want_cs else None,
                                      sy if want_cs else None, M, scale32, scale_kind,
                                      clamp_max, eps_ls)
        ctx.save_for_backward(xhat, yhat, invx, invy, row_lse, col_lse, scale32, logits_scale)
        ctx.cfg = (eps_ls, normalize, clamp_max, scale_kind, impl, in_dtype)
        ctx.kept = (e_keep, e_flag)  # consumed by the (single) backward, never an output
        return loss

    @staticmethod
    def backward(ctx, gout):
        xhat, yhat, invx, invy, row_lse, col_lse, scale32, logits_scale = ctx.saved_tensors
        eps_ls, normalize, clamp_max, scale_kind, impl, in_dtype = ctx.cfg
        B, M, _ = xhat.shape
        n = float(M)
        gout = gout.detach().to(torch.float32).reshape(B).contiguous()
        e_keep, e_flag = ctx.kept
        ctx.kept = (None, None)  # E is overwritten by G': a second backward recomputes
        weights = (0.5 / n, 0.5 / n, (1.0 - eps_ls) / n, eps_ls / (n * n))
        need_da = ctx.needs_input_grad[0] or ctx.needs_input_grad[2]
        bound = ops.DOT_BOUND_NORMALIZED if normalize else 0.0
        if ctx.split is not None:
            x4, y4, xh, xl, yh, yl = ctx.split
            da, db, dscale = ops.sim_bwd_split(x4, y4, xh, xl, yh, yl, scale32, scale_kind,
                                               clamp_max, 0, row_lse, col_lse, *weights, gout,
                                               need_da, ctx.needs_input_grad[1], bound)
        else:
            da, db, dscale, _ = ops.sim_bwd(
                xhat, yhat, None, None, None, scale32, scale_kind, clamp_max, ops.SIM_PLAIN, 0,
                row_lse, col_lse, *weights, gout, need_da, ctx.needs_input_grad[1], False,
                impl, bound, e_keep, e_flag)
        out_dt = ops.BF16 if in_dtype == torch.bfloat16 else ops.F32
        dimg = dtxt = dls = None
        if ctx.needs_input_grad[0]:
            dimg = ops.rownorm_bwd(da, xhat, invx, normalize, out_dt, None).to(in_dtype)
        if ctx.needs_input_grad[1]:
            dtxt = ops.rownorm_bwd(db, yhat, invy, normalize, out_dt, None).to(in_dtype)
        if ctx.needs_input_grad[2]:
            dls = ops.scale_param_grad(dscale, logits_scale, scale_kind, clamp_max)
        return dimg, dtxt, dls, None, None, None, None, None, None


def clip_loss_batched(img, txt, logits_scale, eps_ls, normalize, clamp_max,
                      scale_kind=ops.SCALE_LOG, impl=ops.IMPL_AUTO, compute_bf16=False):
    """[B, M, d] x2 -> loss [B] (fp32)."""
    if not torch.is_tensor(logits_scale):
        logits_scale = torch.tensor(float(logits_scale), device=img.device)
    return _ClipLossFn.apply(img, txt, logits_scale, float(eps_ls), bool(normalize),
                             float(clamp_max), int(scale_kind), int(impl), bool(compute_bf16))


class ClipLoss(nn.Module):
    """
    Simple CLIP contrastive loss for single-GPU training (reference clipLoss.py:13-22).
    Computes contrastive loss between image and text features.
    """

    def __init__(self, label_smoothing: float = 0.1):
        super().__init__()
        self.label_smoothing = label_smoothing
        self.register_buffer("smoothing", torch.tensor(self.label_smoothing))
        self.impl = ops.IMPL_AUTO
        # None: compute in the input dtype (reference semantics).  torch.bfloat16: fp32 inputs
        # run the bf16 tcgen05 kernels (bf16 contract 2e-3 / 2e-2); also taken under
        # torch.autocast(dtype=torch.bfloat16), where the reference's matmul is bf16 as well.
        self.compute_dtype = None

    def _bf16(self):
        return self.compute_dtype == torch.bfloat16 or (
            torch.is_autocast_enabled("cuda") and
            torch.get_autocast_dtype("cuda") == torch.bfloat16)

    def forward_sampled(self, image_features: torch.Tensor, text_features: torch.Tensor,
                        logits_scale: torch.Tensor):
        """Per-slice loss [S] of [S, B, E] samples; always normalises, clamp 3.912
        (reference clipLoss.py:24-79)."""
        B1, B2, E = image_features.shape
        return clip_loss_batched(image_features, text_features, logits_scale,
                                 self.label_smoothing, True, 3.912, impl=self.impl,
                                 compute_bf16=self._bf16())

    def forward(self, image_features: torch.Tensor, text_features: torch.Tensor,
                logits_scale: torch.Tensor, normalize=True):
        """Scalar fp32 loss; clamp 4.6052 (reference clipLoss.py:81-121)."""
        loss = clip_loss_batched(image_features.unsqueeze(0), text_features.unsqueeze(0),
                                 logits_scale, self.label_smoothing, normalize, 4.6052,
                                 impl=self.impl, compute_bf16=self._bf16())
        return loss.reshape(())

"""Gradient accumulation with cached negatives (reference AlignCLIP/main/train.py:119-163).

The reference caches the detached features of `accum_freq` micro-batches, then for micro-batch j
re-runs the model, splices the LIVE features into the cache

    inputs[key] = torch.cat(accumulated[:j] + [model_out[key]] + accumulated[j + 1:])   (:153)

and calls the loss on the N = accum_freq * b concatenated pairs: a full N x N forward whose
gradient reaches only the b live rows of each modality (and `logit_scale`).  `ClipLoss` on the
spliced tensors already does that through autograd; this module does the same arithmetic with a
backward that touches only what has a gradient:

    dX_live = s G[live, :]  Y_all      one rectangular [b, N] recompute + GEMM (diag_offset = j b)
    dY_live = s G[:, live]^T X_all     the same kernels with the roles of rows / columns swapped
    d s     = sum_ij G_ij z_ij / s     needs every row: taken from a full row pass only when
                                        `logit_scale` requires a gradient

i.e. (2 + 8 / accum_freq) N^2 d FLOPs per micro-step with a frozen scale instead of 6 N^2 d.
The loss value is the reference's `ClipLoss.forward` (loss.py:117-132) on the spliced batch.
"""
from __future__ import annotations

from typing import Sequence

import torch

from .. import ops as _ops


class _LiveRowsClipFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, live_img, live_txt, logit_scale, all_img, all_txt, row0, impl, backend):
        B = backend
        dt = all_img.dtype
        x3 = all_img.detach().unsqueeze(0).contiguous()
        y3 = all_txt.detach().unsqueeze(0).contiguous()
        sc = logit_scale.detach().reshape(1).float().contiguous()
        N = x3.shape[1]
        row_lse, cmax, csum, diag = B.sim_fwd(x3, y3, None, None, None, sc, B.SCALE_MUL, 0.0,
                                              B.SIM_PLAIN, 0, True, impl)
        col_lse = B.col_merge(cmax.reshape(1, -1), csum.reshape(1, -1)).reshape(1, -1)
        loss = B.clip_loss_finalize(row_lse, diag, col_lse, None, None, N, sc, B.SCALE_MUL, 0.0,
                                    0.0)
        ctx.save_for_backward(x3, y3, sc, row_lse, col_lse, logit_scale)
        ctx.cfg = (int(row0), int(live_img.shape[0]), impl, B, dt)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, gout):
        x3, y3, sc, row_lse, col_lse, logit_scale = ctx.saved_tensors
        row0, b, impl, B, dt = ctx.cfg
        N = float(x3.shape[1])
        g = gout.detach().float().reshape(1).contiguous()
        w = (0.5 / N, 0.5 / N, 1.0 / N, 0.0)
        live = slice(row0, row0 + b)
        need_img, need_txt, need_scale = ctx.needs_input_grad[:3]
        dimg = dtxt = dls = None
        if need_scale:
            # every row of G enters d/ds: one full row pass (dX of all rows), the live rows of
            # which are the image gradient
            da, _, dscale, _ = B.sim_bwd(x3, y3, None, None, None, sc, B.SCALE_MUL, 0.0,
                                         B.SIM_PLAIN, 0, row_lse, col_lse, *w, g, True, False,
                                         False, impl)
            dls = B.scale_param_grad(dscale, logit_scale, B.SCALE_MUL, 0.0)
            if need_img:
                dimg = da[0, live].to(dt)
        elif need_img:
            da, _, _, _ = B.sim_bwd(x3[:, live].contiguous(), y3, None, None, None, sc,
                                    B.SCALE_MUL, 0.0, B.SIM_PLAIN, row0,
                                    row_lse[:, live].contiguous(), col_lse, *w, g, True, False,
                                    False, impl)
            dimg = da[0].to(dt)
        if need_txt:
            # rows = live texts, columns = all images: row statistics are the column LSEs
            db, _, _, _ = B.sim_bwd(y3[:, live].contiguous(), x3, None, None, None, sc,
                                    B.SCALE_MUL, 0.0, B.SIM_PLAIN, row0,
                                    col_lse[:, live].contiguous(), row_lse, *w, g, True, False,
                                    False, impl)
            dtxt = db[0].to(dt)
        return dimg, dtxt, dls, None, None, None, None, None


def cached_negatives_clip_loss(image_features: torch.Tensor, text_features: torch.Tensor,
                               logit_scale, accum_image: Sequence[torch.Tensor],
                               accum_text: Sequence[torch.Tensor], j: int, output_dict=False,
                               impl: int = _ops.IMPL_AUTO, backend=None):
    """open_clip `ClipLoss` (loss.py:117-132, world_size 1) of micro-step j of the reference's
    accumulation loop (main/train.py:139-157): `image_features` / `text_features` are the live
    outputs of micro-batch j, `accum_image` / `accum_text` the lists of cached (no-grad)
    features of all `accum_freq` micro-batches (entry j is ignored, as in :153).  Returns the
    loss of the spliced batch; gradients reach the live features and `logit_scale` only."""
    B = backend if backend is not None else _ops
    if not torch.is_tensor(logit_scale):
        logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
    if not (0 <= j < len(accum_image)) or len(accum_image) != len(accum_text):
        raise IndexError("micro-batch index %d outside the %d cached batches" % (j, len(accum_image)))
    row0 = sum(int(t.shape[0]) for t in accum_image[:j])
    if row0 != sum(int(t.shape[0]) for t in accum_text[:j]):
        raise ValueError("cached image / text batches must have matching sizes")
    dt = image_features.dtype
    all_img = torch.cat([t.detach().to(dt) for t in accum_image[:j]] + [image_features.detach()] +
                        [t.detach().to(dt) for t in accum_image[j + 1:]])
    all_txt = torch.cat([t.detach().to(dt) for t in accum_text[:j]] + [text_features.detach()] +
                        [t.detach().to(dt) for t in accum_text[j + 1:]])
    loss = _LiveRowsClipFn.apply(image_features, text_features, logit_scale, all_img, all_txt,
                                 row0, int(impl), B)
    return {"contrastive_loss": loss} if output_dict else loss

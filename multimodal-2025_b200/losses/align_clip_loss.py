"""Drop-in for the loss classes of the reference's `AlignCLIP/align_clip/loss.py`
(`ClipLoss` :67-132 and `ClipInModalityLoss` :135-273), single-process branch.

`logit_scale` is ALREADY exponentiated here (open_clip convention).  Nothing N x N is
materialised: every cross-entropy is `mean_i (LSE_j z_ij - z_ii)` from the fused kernels.

  cross-modal   alpha * [CE(s I T^T) + CE(s T I^T)]                      (loss.py:267-269, a sum)
  in-modality   beta  * [CE(s I I^T) + CE(s T T^T)]                      (loss.py:258-265)
  semantic      beta  * CE(s [(T T^T) o (1 - S S^T) + diag(i_i . t_i)])   (loss.py:229-255)

The semantic term runs the `MODULATED` mode of mm25_sim_fwd / mm25_sim_bwd: a second accumulator
for the normalised semantic features and the paired dot products added on the diagonal.
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops
from .clipLoss import clip_loss_batched


def _as_scale(logit_scale, ref):
    if not torch.is_tensor(logit_scale):
        logit_scale = torch.tensor(float(logit_scale), device=ref.device)
    return logit_scale


def _ce_pair(a, b, logit_scale):
    """0.5 * (CE(s a b^T) + CE(s b a^T)), labels = arange: the CLIP kernels with eps = 0."""
    return clip_loss_batched(a.unsqueeze(0), b.unsqueeze(0), logit_scale, 0.0, False, 0.0,
                             scale_kind=ops.SCALE_MUL).reshape(())


class _ModulatedCEFn(torch.autograd.Function):
    """mean_i (LSE_j z_ij - z_ii),  z = s [(f f^T) o (1 - u u^T) + diag(D)]  (loss.py:229-242)."""

    @staticmethod
    def forward(ctx, feats, sem_hat, paired, logit_scale, impl=ops.IMPL_AUTO):
        in_dtype = feats.dtype
        f = feats.detach()
        if f.dtype not in (torch.float32, torch.bfloat16):
            f = f.float()
        f3 = f.contiguous().unsqueeze(0)
        u3 = sem_hat.detach().float().contiguous().unsqueeze(0)
        D = paired.detach().float().contiguous().reshape(1, -1)
        sc = logit_scale.detach().reshape(1).float().contiguous()
        n = f3.shape[1]
        # bf16 features of an eligible shape run the two-accumulator tcgen05 kernels
        # (csrc/sim_tc_mod.cu); everything else the fp32 SIMT kernels (include/mm25.h)
        row_lse, _, _, diag = ops.sim_fwd(f3, f3, u3, u3, D, sc, ops.SCALE_MUL, 0.0,
                                          ops.SIM_MODULATED, 0, False, impl)
        # (1/n) sum_i (r_i - z_ii): the CLIP finalize with weights 1/(2n) on 2 r_i and 1/n on diag
        loss = ops.clip_loss_finalize(row_lse * 2.0, diag, None, None, None, n, sc,
                                      ops.SCALE_MUL, 0.0, 0.0)
        ctx.save_for_backward(f3, u3, D, sc, row_lse, logit_scale)
        ctx.in_dtype = in_dtype
        ctx.impl = impl
        return loss.reshape(())

    @staticmethod
    def backward(ctx, gout):
        f3, u3, D, sc, row_lse, logit_scale = ctx.saved_tensors
        n = float(f3.shape[1])
        g = gout.detach().float().reshape(1).contiguous()
        da, db, dscale, ddiag = ops.sim_bwd(f3, f3, u3, u3, D, sc, ops.SCALE_MUL, 0.0,
                                            ops.SIM_MODULATED, 0, row_lse, None, 1.0 / n, 0.0,
                                            1.0 / n, 0.0, g, True, True, True, ctx.impl)
        dfeats = (da[0] + db[0]).to(ctx.in_dtype) if ctx.needs_input_grad[0] else None
        dpaired = ddiag[0] if ctx.needs_input_grad[2] else None
        dls = None
        if ctx.needs_input_grad[3]:
            dls = ops.scale_param_grad(dscale, logit_scale, ops.SCALE_MUL, 0.0)
        return dfeats, None, dpaired, dls, None


def _sharded_args(mod):
    """Map the reference's multi-process flags (loss.py:20-64, 103-114) onto the sharded
    modules.  With `local_loss=False` the reference evaluates the GLOBAL loss on every rank:
    `gather_with_grad=False` lets only the rank's own slot of the gathered features carry a
    gradient (loss.py:55-58) -> d(global loss)/d(local shard), `grad_scale = 1`;
    `gather_with_grad=True` back-propagates through `torch.distributed.nn.all_gather`, which
    SUMS the P identical per-rank losses' gradients -> `grad_scale = world_size` (the open_clip
    convention; DDP's gradient averaging then undoes the factor).  `local_loss=True` (each rank's
    own loss over its local rows, loss.py:108-110) is built for `ClipLoss`
    (dist.sharded_local_clip_loss) and raises for `ClipInModalityLoss`; Horovod raises."""
    import torch.distributed as dist
    if mod.use_horovod:
        raise NotImplementedError("use_horovod=True: Horovod collectives are out of scope "
                                  "(torch.distributed only)")
    if mod.local_loss and not getattr(mod, "_local_loss_ok", False):
        raise NotImplementedError(
            "local_loss=True (per-rank loss over the local rows, loss.py:191-200) is not "
            "implemented for ClipInModalityLoss; use local_loss=False (global loss, identical "
            "on every rank)")
    group = getattr(mod, "group", None)
    if not dist.is_available() or not dist.is_initialized():
        raise RuntimeError("world_size > 1 needs an initialised torch.distributed process group")
    g = group if group is not None else dist.group.WORLD
    ws, rk = dist.get_world_size(g), dist.get_rank(g)
    if ws != int(mod.world_size) or rk != int(mod.rank):
        raise ValueError("rank/world_size (%d/%d) do not match the process group (%d/%d)"
                         % (mod.rank, mod.world_size, rk, ws))
    return group, float(ws) if mod.gather_with_grad else 1.0


class ClipLoss(nn.Module):
    """open_clip-style CLIP loss (reference AlignCLIP/align_clip/loss.py:67-132); world_size > 1
    delegates to dist.ShardedAlignClipLoss (needs an initialised process group)."""

    def __init__(self, local_loss=False, gather_with_grad=False, cache_labels=False, rank=0,
                 world_size=1, use_horovod=False):
        super().__init__()
        self.local_loss = local_loss
        self.gather_with_grad = gather_with_grad
        self.cache_labels = cache_labels
        self.rank = rank
        self.world_size = world_size
        self.use_horovod = use_horovod
        self.group = None    # optional process group (attribute, not in the reference signature)
        self._backend = None  # tests inject the CPU stand-in ops (gloo)

    def forward(self, image_features, text_features, logit_scale, output_dict=False,
                semantic_features=None):
        if self.world_size > 1:
            # global-batch loss from the local shards (local_loss=False, loss.py:103-114);
            # gradient convention per gather_with_grad, see _sharded_args
            from ..dist import ShardedAlignClipLoss, sharded_local_clip_loss
            self._local_loss_ok = True
            group, grad_scale = _sharded_args(self)
            if self.local_loss:   # this rank's own loss over its local rows (loss.py:108-110)
                loss = sharded_local_clip_loss(image_features, text_features, logit_scale,
                                               group, self._backend, self.gather_with_grad)
                return {"contrastive_loss": loss} if output_dict else loss
            crit = ShardedAlignClipLoss(group, grad_scale, self._backend)
            return crit(image_features, text_features, logit_scale, output_dict=output_dict)
        logit_scale = _as_scale(logit_scale, image_features)
        clip_loss = _ce_pair(image_features, text_features, logit_scale)
        return {"contrastive_loss": clip_loss} if output_dict else clip_loss


class ClipInModalityLoss(nn.Module):
    """reference AlignCLIP/align_clip/loss.py:135-273; world_size > 1 delegates to
    dist.ShardedClipInModalityLoss."""

    def __init__(self, local_loss=False, gather_with_grad=False, cache_labels=False, rank=0,
                 world_size=1, use_horovod=False, alpha=1.0, beta=0.5, n_epoch=30,
                 nl_semantic_supervision=False, separate_text=True, separate_image=False):
        super().__init__()
        self.local_loss = local_loss
        self.gather_with_grad = gather_with_grad
        self.cache_labels = cache_labels
        self.rank = rank
        self.world_size = world_size
        self.use_horovod = use_horovod
        self.alpha = alpha
        self.beta = beta
        self.nl_semantic_supervision = nl_semantic_supervision
        self.separate_text = separate_text
        self.separate_image = separate_image
        self.group = None    # optional process group (attribute, not in the reference signature)
        self._backend = None  # tests inject the CPU stand-in ops (gloo)

    def forward(self, image_features, text_features, logit_scale, output_dict=False,
                semantic_features=None):
        if self.world_size > 1:
            # the reference's own branch never gathers the semantic features; the
            # sharded module evaluates the single-process formula on the concatenated batch
            from ..dist import ShardedClipInModalityLoss
            group, grad_scale = _sharded_args(self)
            crit = ShardedClipInModalityLoss(self.alpha, self.beta, self.nl_semantic_supervision,
                                             self.separate_text, self.separate_image,
                                             group=group, grad_scale=grad_scale,
                                             backend=self._backend)
            return crit(image_features, text_features, logit_scale, output_dict=output_dict,
                        semantic_features=semantic_features)
        logit_scale = _as_scale(logit_scale, image_features)
        if self.nl_semantic_supervision:
            if semantic_features is None:
                raise TypeError("nl_semantic_supervision=True needs semantic_features")
            sem = semantic_features.detach().float().unsqueeze(0)
            # semantic_features / ||.|| with no eps (loss.py:229)
            sem_hat, _, _ = ops.rownorm_fwd(sem, True, 0.0, ops.F32, False)
            # diag(I T^T): the paired similarities added back on the diagonal (loss.py:237)
            paired = (image_features.float() * text_features.float()).sum(dim=-1)
            terms = []
            if self.separate_text:
                terms.append(_ModulatedCEFn.apply(text_features, sem_hat[0], paired, logit_scale))
            if self.separate_image:
                terms.append(_ModulatedCEFn.apply(image_features, sem_hat[0], paired,
                                                  logit_scale))
            if not terms:
                # the reference leaves inModality_loss undefined here (UnboundLocalError)
                raise UnboundLocalError("inModality_loss: neither separate_text nor "
                                        "separate_image is set (reference loss.py:249-255)")
            inModality_loss = self.beta * sum(terms[1:], terms[0])
        else:
            inModality_loss = self.beta * (
                _ce_pair(image_features, image_features, logit_scale) +
                _ce_pair(text_features, text_features, logit_scale))
        clip_loss = self.alpha * (2.0 * _ce_pair(image_features, text_features, logit_scale))
        total_loss = inModality_loss + clip_loss
        if output_dict:
            return {"total_loss": total_loss, "clip_loss": clip_loss,
                    "inModality_loss": inModality_loss}
        return total_loss

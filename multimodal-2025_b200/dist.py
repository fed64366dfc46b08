"""Row-sharded CLIP loss across the GPUs of one NVSwitch box (one process per GPU, NCCL).

The reference has no working multi-process path; its semantic precedent is
`gather_features(..., gather_with_grad=True)` + `local_loss=False` in
AlignCLIP/align_clip/loss.py:20-64,103-114.  Algorithm:

  forward   rank r owns rows [r n, (r+1) n) of both modalities (n = N / P)
            1. local row-normalise; all-reduce the [d] column sums (label-smoothing term)
            2. all-gather Yhat (bf16/fp32, N d elements in total)
            3. fused kernel on the local block [n, N]: complete row LSE, per-column partial
               (max, sum) over the local rows, positives at column offset r n (loss.py:94-95)
            4. all-gather the [N] column partials, merge in fixed rank order (deterministic)
            5. all-reduce the scalar partial losses
  backward  recompute the local block; dXhat_local is complete locally; the partial
            dYhat_all [N, d] is reduce-scattered (sum) so each rank receives its own rows;
            d(scale) is all-reduced.

The returned loss is the GLOBAL loss (identical on every rank) and the gradients are those of
the global loss w.r.t. the local shards (times the upstream gradient, assumed equal on all
ranks) — what `gather_with_grad=True, local_loss=False` computes divided by P.  `grad_scale`
multiplies them (set it to P to reproduce open_clip's sum-over-ranks convention).

Kernels are reached through `backend` (default: multimodal_2025_b200.ops, CUDA only); tests
inject a CPU stand-in to exercise the collective logic under gloo.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
from torch import nn

_EPS_NORM = 1e-12


def _all_gather_cat(t: torch.Tensor, group) -> torch.Tensor:
    """[n, ...] per rank -> [P n, ...] (rank order)."""
    world = dist.get_world_size(group)
    t = t.contiguous()
    out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    if dist.get_backend(group) == "gloo":
        parts = list(out.chunk(world, dim=0))
        dist.all_gather(parts, t, group=group)
    else:
        dist.all_gather_into_tensor(out, t, group=group)
    return out


def _reduce_scatter_sum(t: torch.Tensor, group) -> torch.Tensor:
    """[P n, ...] per rank -> sum over ranks of the local [n, ...] slice."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    t = t.contiguous()
    n = t.shape[0] // world
    if dist.get_backend(group) == "gloo":  # gloo has no reduce_scatter
        dist.all_reduce(t, group=group)
        return t[rank * n:(rank + 1) * n].clone()
    out = torch.empty((n,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.reduce_scatter_tensor(out, t, group=group)
    return out


_SYMM_CACHE = {}
LAST_BACKWARD_PATH = None  # "fused gemm+reduce-scatter (peer memory)" or "nccl reduce-scatter"


def _agree(ok: bool, device, group) -> bool:
    """All ranks take the symmetric-memory fast path or none does: a rendezvous that failed on
    one rank only would leave that rank in NCCL collectives while the others wait in peer
    barriers.  One MIN all-reduce per buffer creation (results are cached)."""
    try:
        flag = torch.tensor([1 if ok else 0], device=device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        return bool(int(flag.item()))
    except Exception:  # noqa: BLE001 - no usable collective: stay on the conservative path
        return False


def _group_key(group):
    """Cache key of a process group that survives garbage collection of the Python object
    (id() values are reused): its global ranks."""
    try:
        return tuple(dist.get_process_group_ranks(group))
    except Exception:  # noqa: BLE001
        return id(group)


def release_symmetric_buffers():
    """Drop every cached symmetric-memory buffer (they are created once per shape and group and
    otherwise live until exit)."""
    _SYMM_CACHE.clear()
    _SYMM_GEN.clear()
    _GRAD_CLEAN.clear()


def _symm_grad_buffer(n: int, d: int, device, group):
    """[n, d] fp32 buffer in symmetric memory (peer-mapped on every rank of `group`) that the
    fused backward scatter-adds dY into.  Returns (tensor, handle) or None when symmetric
    memory cannot be set up (then the NCCL reduce-scatter is used)."""
    key = (n, d, device.index, _group_key(group))
    if key in _SYMM_CACHE:
        return _SYMM_CACHE[key]
    out = None
    try:
        import torch.distributed._symmetric_memory as symm_mem
        buf = symm_mem.empty((n, d), dtype=torch.float32, device=device)
        hdl = symm_mem.rendezvous(buf, group)
        out = (buf, hdl)
    except Exception:  # no NVLink peer access / unsupported allocator: keep the NCCL path
        out = None
    if not _agree(out is not None, device, group):
        out = None
    _SYMM_CACHE[key] = out
    return out


_SYMM_GEN = {}
_GRAD_CLEAN = {}  # symmetric dY buffer zeroed (and published) by the forward of the same step


def _symm_buffer(tag, shape, dtype, device, group):
    """Cached symmetric-memory tensor + handle (peer-mapped on every rank), or None."""
    key = (tag, tuple(shape), dtype, device.index, _group_key(group))
    if key in _SYMM_CACHE:
        return key, _SYMM_CACHE[key]
    out = None
    try:
        import torch.distributed._symmetric_memory as symm_mem
        buf = symm_mem.empty(tuple(shape), dtype=dtype, device=device)
        hdl = symm_mem.rendezvous(buf, group)
        out = (buf, hdl)
    except Exception:  # no peer access / unsupported allocator: callers keep the NCCL path
        out = None
    if not _agree(out is not None, device, group):
        out = None
    _SYMM_CACHE[key] = out
    return key, out


def _peer_all_gather(B, t, tag, group, barrier=True):
    """[n, ...] per rank -> [P n, ...]: every rank stores its shard straight into all peers'
    symmetric buffers over NVLink (one kernel) and one symmetric-memory barrier follows — no NCCL
    ring, no staging.  Returns (gathered buffer, key) or None when the fast path does not apply.
    The buffer is reused by the next call with the same shape: consume it (or check
    `_SYMM_GEN[key]`) before gathering again."""
    if os.environ.get("MM25_DIST_LEGACY") and not barrier:
        return None
    if not (hasattr(B, "peer_bcast") and t.is_cuda and dist.get_backend(group) == "nccl"):
        return None
    nbytes = t.numel() * t.element_size()
    if nbytes % 4 or os.environ.get("MM25_NCCL_GATHER"):
        return None
    if nbytes % 16 and nbytes > (1 << 20):
        return None   # mm25_peer_bcast moves large payloads as 16-byte vectors only: NCCL path
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    key, sb = _symm_buffer(tag, (world * t.shape[0],) + tuple(t.shape[1:]), t.dtype, t.device,
                           group)
    if sb is None:
        return None
    buf, hdl = sb
    mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
    if (mc and nbytes % 16 == 0 and nbytes >= (1 << 16) and hasattr(B, "peer_bcast_mc")
            and os.environ.get("MM25_MC", "1") == "1"):
        # NVSwitch multicast: the shard is sent once and replicated in the switch
        B.peer_bcast_mc(t, mc, rank * nbytes)
    else:
        B.peer_bcast(t, [int(p) for p in hdl.buffer_ptrs], rank * nbytes)
    if barrier:  # barrier=False: the caller's next symmetric-memory barrier publishes the stores
        hdl.barrier(channel=0)
    _SYMM_GEN[key] = _SYMM_GEN.get(key, 0) + 1
    return buf, key


def _all_reduce_scalar(B, t, tag, group):
    """Sum of a [1] fp32 tensor over the ranks: peer stores into a [P, 1] symmetric buffer, one
    barrier and a local fixed-order sum (same bits on every rank); NCCL all-reduce otherwise."""
    pg = _peer_all_gather(B, t.reshape(1, 1), tag, group)
    if pg is None:
        dist.all_reduce(t, group=group)
        return t
    return pg[0].sum(dim=0)


class _ShardedClipFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, txt, logits_scale, eps_ls, normalize, clamp_max, group, backend,
                grad_scale, impl, scale_kind=1):
        B = backend
        world = dist.get_world_size(group)
        rank = dist.get_rank(group)
        in_dtype = img.dtype
        if in_dtype not in (torch.float32, torch.bfloat16):
            img, txt = img.float(), txt.float()
        if txt.dtype != img.dtype:
            txt = txt.to(img.dtype)
        cdt = B.BF16 if img.dtype == torch.bfloat16 else B.F32
        n, d = img.shape
        N = n * world
        want_cs = eps_ls != 0.0
        xhat, invx, sx = B.rownorm_fwd(img.unsqueeze(0), normalize, _EPS_NORM, cdt, want_cs)
        yhat, invy, sy = B.rownorm_fwd(txt.unsqueeze(0), normalize, _EPS_NORM, cdt, want_cs)
        scale32 = logits_scale.detach().reshape(1).to(torch.float32).contiguous()
        # (2) all-gather the other modality
        pg = _peer_all_gather(B, yhat[0], "yall", group)
        if pg is not None:
            yall, ykey = pg[0].unsqueeze(0), pg[1]
            ygen = _SYMM_GEN[ykey]
        else:
            yall, ykey, ygen = _all_gather_cat(yhat[0], group).unsqueeze(0), None, 0
        # (3) local block
        bound = 1.01 if normalize else 0.0
        keep = (hasattr(B, "sim_fwd_keep") and xhat.is_cuda
                and any(ctx.needs_input_grad[:3]) and bound > 0.0 and impl != 1
                and B.tc_eligible_bwd(xhat, yall) and 2 * n * N <= B.keep_e_bytes_limit(n, N, xhat.shape[-1]))
        e_keep = e_flag = None
        # fp32 shards of an eligible shape keep fp32-grade arithmetic on the tensor cores (two-term
        # bf16 splits, include/mm25.h mm25_split_bf16x2) instead of the FFMA kernels
        split = (cdt == B.F32 and impl == 0 and xhat.is_cuda and hasattr(B, "split_bf16x2")
                 and os.environ.get("MM25_FP32_TC", "1") != "0"
                 and B.sim_split_eligible(1, n, N, xhat.shape[-1]))
        ctx.split = None
        if split:
            x4, xh, xl = B.split_bf16x2(xhat, 0)
            y4, yh, yl = B.split_bf16x2(yall, 1)
            row_lse, cmax, csum, diag = B.sim_fwd(x4, y4, None, None, None, scale32, scale_kind,
                                                  clamp_max, B.SIM_PLAIN, rank * n, True, impl,
                                                  bound)
            ctx.split = (x4, y4, xh, xl, yh, yl)
        elif keep:  # E = exp(z - shift) of the local [n, N] block is kept for the backward
            row_lse, cmax, csum, diag, e_keep, e_flag = B.sim_fwd_keep(
                xhat, yall, scale32, scale_kind, clamp_max, rank * n, True, impl, bound)
        else:
            row_lse, cmax, csum, diag = B.sim_fwd(xhat, yall, None, None, None, scale32,
                                                  scale_kind, clamp_max, B.SIM_PLAIN, rank * n,
                                                  True, impl, bound)
        # (4) ONE all-gather carries the column partials (max, sum)[N], each rank's column sums
        #     of xhat / yhat (smoothing term) and the scalar loss piece of its own ROWS;
        #     everything is then merged / summed locally in rank order (deterministic).  The
        #     column part of the loss needs no exchange at all: after the merge every rank
        #     holds all N column LSEs and sums them itself, so the forward has no scalar
        #     all-reduce (one kernel + one barrier fewer than round 1).
        legacy = bool(os.environ.get("MM25_DIST_LEGACY"))
        extra = torch.cat([sx.reshape(-1), sy.reshape(-1)]) if want_cs else cmax.new_zeros(0)
        if legacy:
            payload = torch.cat([cmax[0], csum[0], extra]).unsqueeze(0)
        else:
            rowpart = B.clip_loss_finalize(row_lse, diag, None, None, None, N, scale32,
                                           scale_kind, clamp_max, eps_ls)
            payload = torch.cat([cmax[0], csum[0], extra, rowpart.reshape(1),
                                 cmax.new_zeros(3)]).unsqueeze(0)   # 16-byte multiple
        # the fused backward scatter-adds dY into this rank's symmetric buffer: zero it HERE so
        # that the exchange below publishes the zeroing (the backward then starts without its
        # own zero + barrier); _GRAD_CLEAN tells the backward whether that still holds
        gkey = None
        if (not legacy and any(ctx.needs_input_grad[:3]) and hasattr(B, "sim_bwd_rs")
                and xhat.is_cuda and xhat.dtype == torch.bfloat16
                and dist.get_backend(group) == "nccl" and B.tc_eligible_bwd(xhat, yall)):
            symm = _symm_grad_buffer(n, xhat.shape[-1], xhat.device, group)
            if symm is not None:
                symm[0].zero_()
                gkey = (n, xhat.shape[-1], xhat.device.index, _group_key(group))
                _GRAD_CLEAN[gkey] = True
        pg = _peer_all_gather(B, payload, "colstats", group)            # [P, 2N (+ 2d) (+ 4)]
        gathered = pg[0] if pg is not None else _all_gather_cat(payload, group)
        if hasattr(B, "col_merge_ld"):
            col_lse = B.col_merge_ld(gathered, N)
        else:
            col_lse = B.col_merge(gathered[:, :N].contiguous(), gathered[:, N:2 * N].contiguous())
        if legacy:
            # (5) scalar pieces of the loss: local rows + local columns, then a small all-reduce
            part = B.clip_loss_finalize(row_lse, diag, col_lse[rank * n:(rank + 1) * n]
                                        .reshape(1, n).contiguous(), None, None, N, scale32,
                                        scale_kind, clamp_max, eps_ls)
            red = _all_reduce_scalar(B, part.reshape(1), "loss", group)
        else:
            off = 2 * N + (2 * d if want_cs else 0)
            if hasattr(B, "sharded_loss") and gathered.is_contiguous():
                red = B.sharded_loss(gathered, off, col_lse, 2 * N, 2 * N + d,
                                     d if want_cs else 0, scale32, scale_kind, clamp_max,
                                     eps_ls / float(N) ** 2)
                want_cs = False   # the smoothing term is inside the fused tail
            else:
                red = (gathered[:, off].sum(dtype=torch.float64) +
                       col_lse.sum(dtype=torch.float64) * (0.5 / N)).to(torch.float32).reshape(1)
        if want_cs:
            coef = eps_ls / float(N) ** 2
            if hasattr(B, "smooth_term_"):
                red = B.smooth_term_(red, gathered, 2 * N, 2 * N + d, d, scale32, scale_kind,
                                     clamp_max, coef)
            else:
                sxg = gathered[:, 2 * N:2 * N + d].sum(dim=0)
                syg = gathered[:, 2 * N + d:2 * N + 2 * d].sum(dim=0)
                s = (scale32.clamp(max=clamp_max).exp() if scale_kind == B.SCALE_LOG
                     else scale32).reshape(())
                red = red - coef * s * (sxg * syg).sum()
        loss = red.reshape(())
        ctx.save_for_backward(xhat, yhat, yall, invx, invy, row_lse, col_lse, scale32,
                              logits_scale)
        ctx.cfg = (eps_ls, normalize, clamp_max, group, backend, grad_scale, impl, in_dtype,
                   rank, n, N, scale_kind)
        ctx.kept = (e_keep, e_flag)
        ctx.ygen = (ykey, ygen)
        ctx.gkey = gkey
        return loss

    @staticmethod
    def backward(ctx, gout):
        (xhat, yhat, yall, invx, invy, row_lse, col_lse, scale32,
         logits_scale) = ctx.saved_tensors
        ykey, ygen = ctx.ygen
        if ykey is not None and _SYMM_GEN.get(ykey) != ygen:
            # another forward reused the symmetric gather buffer since: gather Yhat again
            yall = _all_gather_cat(yhat[0], ctx.cfg[3]).unsqueeze(0)
        (eps_ls, normalize, clamp_max, group, B, grad_scale, impl, in_dtype, rank, n,
         N, scale_kind) = ctx.cfg
        nn_ = float(N)
        g = gout.detach().to(torch.float32).reshape(1)
        if grad_scale != 1.0:
            g = g * grad_scale
        g = g.contiguous()
        bound = 1.01 if normalize else 0.0
        e_keep, e_flag = ctx.kept
        ctx.kept = (None, None)
        kept = () if e_keep is None else (e_keep, e_flag)
        weights = (0.5 / nn_, 0.5 / nn_, (1.0 - eps_ls) / nn_, eps_ls / (nn_ * nn_))
        symm = None
        if (hasattr(B, "sim_bwd_rs") and xhat.is_cuda and xhat.dtype == torch.bfloat16 and
                dist.get_backend(group) == "nccl" and B.tc_eligible_bwd(xhat, yall)):
            symm = _symm_grad_buffer(n, xhat.shape[-1], xhat.device, group)
        pgd = None
        global LAST_BACKWARD_PATH
        LAST_BACKWARD_PATH = ("fused gemm+reduce-scatter (NVLink peer memory)" if symm is not None
                              else "nccl reduce-scatter")
        if ctx.split is not None:
            x4, y4, xh, xl, yh, yl = ctx.split
            da, db, dscale = B.sim_bwd_split(x4, y4, xh, xl, yh, yl, scale32, scale_kind,
                                             clamp_max, rank * n, row_lse, col_lse.reshape(1, N),
                                             *weights, g, True, True, bound)
            dy_local = _reduce_scatter_sum(db[0], group).unsqueeze(0)
            LAST_BACKWARD_PATH = "fp32 split kernels + nccl reduce-scatter"
        elif symm is not None:
            # GEMM -> reduce-scatter in ONE kernel: the dB tiles are added straight into the
            # owner rank's buffer over NVLink peer memory (no NCCL reduce-scatter, no [N, d]
            # fp32 partial in HBM)
            buf, hdl = symm
            if ctx.gkey is not None and _GRAD_CLEAN.get(ctx.gkey):
                _GRAD_CLEAN[ctx.gkey] = False   # zeroed and published by this step's forward
            else:
                buf.zero_()
                hdl.barrier(channel=0)
            da, dscale = B.sim_bwd_rs(xhat, yall, scale32, scale_kind, clamp_max, rank * n,
                                      row_lse, col_lse.reshape(1, N), *weights, g,
                                      [int(p) for p in hdl.buffer_ptrs], bound, *kept)
            if ctx.needs_input_grad[2]:
                # the d/dscale partial travels before the barrier that completes the scatter
                pgd = _peer_all_gather(B, dscale.reshape(1, 1), "dscale", group, barrier=False)
            hdl.barrier(channel=1)
            dy_local = buf.unsqueeze(0)
        else:
            da, db, dscale, _ = B.sim_bwd(xhat, yall, None, None, None, scale32, scale_kind,
                                          clamp_max, B.SIM_PLAIN, rank * n, row_lse,
                                          col_lse.reshape(1, N), *weights, g, True, True, False,
                                          impl, bound, *kept)
            dy_local = _reduce_scatter_sum(db[0], group).unsqueeze(0)
        out_dt = B.BF16 if in_dtype == torch.bfloat16 else B.F32
        dimg = B.rownorm_bwd(da, xhat, invx, normalize, out_dt, None)[0].to(in_dtype)
        dtxt = B.rownorm_bwd(dy_local, yhat, invy, normalize, out_dt, None)[0].to(in_dtype)
        dls = None
        if ctx.needs_input_grad[2]:
            if pgd is not None:
                dscale = pgd[0].sum(dim=0)
            else:
                dscale = _all_reduce_scalar(B, dscale.reshape(1), "dscale", group)
            dls = B.scale_param_grad(dscale, logits_scale, scale_kind, clamp_max)
        return dimg, dtxt, dls, None, None, None, None, None, None, None, None


class ShardedClipLoss(nn.Module):
    """ClipLoss (reference losses/clipLoss.py:81-121) with the batch row-sharded over a process
    group: every rank passes its LOCAL [n, d] image and text shards (same n on all ranks) and
    receives the loss of the GLOBAL batch of P n pairs."""

    def __init__(self, label_smoothing: float = 0.1, group=None, grad_scale: float = 1.0,
                 backend=None):
        super().__init__()
        self.label_smoothing = label_smoothing
        self.register_buffer("smoothing", torch.tensor(self.label_smoothing))
        self.group = group
        self.grad_scale = grad_scale
        self.backend = backend
        self.impl = 0

    def forward(self, image_features, text_features, logits_scale, normalize=True):
        backend = self.backend
        if backend is None:
            from . import ops as backend
        group = self.group if self.group is not None else dist.group.WORLD
        if not torch.is_tensor(logits_scale):
            logits_scale = torch.tensor(float(logits_scale), device=image_features.device)
        return _ShardedClipFn.apply(image_features, text_features, logits_scale,
                                    float(self.label_smoothing), bool(normalize), 4.6052, group,
                                    backend, float(self.grad_scale), int(self.impl))

    def forward_sampled(self, image_features, text_features, logits_scale):
        """ClipLoss.forward_sampled (clipLoss.py:24-79) of the GLOBAL batch: local [S, n, E]
        sample shards -> [S] losses; always normalises, clamp 3.912."""
        backend = self.backend
        if backend is None:
            from . import ops as backend
        group = self.group if self.group is not None else dist.group.WORLD
        if not torch.is_tensor(logits_scale):
            logits_scale = torch.tensor(float(logits_scale), device=image_features.device)
        return _ShardedClipSampledFn.apply(image_features, text_features, logits_scale,
                                           float(self.label_smoothing), 3.912, group, backend,
                                           float(self.grad_scale), int(self.impl))


# ----------------------------------------------------------------------------------------------
# AlignCLIP (reference AlignCLIP/align_clip/loss.py) on the row-sharded global batch
# ----------------------------------------------------------------------------------------------
def _sharded_ce_pair(a, b, logit_scale, group, backend, grad_scale, impl):
    """0.5 (CE(s A B^T) + CE(s B A^T)) of the GLOBAL batch from the local [n, d] shards
    (no smoothing, no normalisation, multiplicative scale: loss.py:103-132)."""
    return _ShardedClipFn.apply(a, b, logit_scale, 0.0, False, 0.0, group, backend,
                                float(grad_scale), int(impl), backend.SCALE_MUL)


class _ShardedLocalCEFn(torch.autograd.Function):
    """open_clip `local_loss=True` (loss.py:108-110, 94-95): rank r's OWN loss
        0.5 [CE(s X_r Y_all^T, labels + r n) + CE(s Y_r X_all^T, labels + r n)]
    from two rectangular [n, N] sweeps (row statistics only, positives at column offset r n).
    `gather_with_grad=True`: the gathered operands carry gradient, i.e. the partial dX_all /
    dY_all of every rank's loss are summed into their owners (reduce-scatter), as
    torch.distributed.nn.all_gather back-propagates; False: only the row operands do."""

    @staticmethod
    def forward(ctx, img, txt, logit_scale, group, backend, gather_with_grad, impl):
        B = backend
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        in_dtype = img.dtype
        x, y = img.detach(), txt.detach()
        if x.dtype not in (torch.float32, torch.bfloat16):
            x, y = x.float(), y.float()
        if y.dtype != x.dtype:
            y = y.to(x.dtype)
        x, y = x.contiguous(), y.contiguous()
        n = x.shape[0]
        xall = _all_gather_cat(x, group).unsqueeze(0)
        yall = _all_gather_cat(y, group).unsqueeze(0)
        sc = logit_scale.detach().reshape(1).float().contiguous()
        x3, y3 = x.unsqueeze(0), y.unsqueeze(0)
        r1, _, _, d1 = B.sim_fwd(x3, yall, None, None, None, sc, B.SCALE_MUL, 0.0, B.SIM_PLAIN,
                                 rank * n, False, impl)
        r2, _, _, d2 = B.sim_fwd(y3, xall, None, None, None, sc, B.SCALE_MUL, 0.0, B.SIM_PLAIN,
                                 rank * n, False, impl)
        # mean_i (r_i - z_ii) per direction: the CLIP finalize with weights 1/(2n) on 2 r_i
        l1 = B.clip_loss_finalize(r1 * 2.0, d1, None, None, None, n, sc, B.SCALE_MUL, 0.0, 0.0)
        l2 = B.clip_loss_finalize(r2 * 2.0, d2, None, None, None, n, sc, B.SCALE_MUL, 0.0, 0.0)
        ctx.save_for_backward(x3, y3, xall, yall, sc, r1, r2, logit_scale)
        ctx.cfg = (group, B, bool(gather_with_grad), impl, in_dtype, rank, n)
        return (0.5 * (l1 + l2)).reshape(())

    @staticmethod
    def backward(ctx, gout):
        x3, y3, xall, yall, sc, r1, r2, logit_scale = ctx.saved_tensors
        group, B, gwg, impl, in_dtype, rank, n = ctx.cfg
        g = (gout.detach().float().reshape(1) * 0.5).contiguous()
        w = 1.0 / float(n)
        dx1, dyall, ds1, _ = B.sim_bwd(x3, yall, None, None, None, sc, B.SCALE_MUL, 0.0,
                                       B.SIM_PLAIN, rank * n, r1, None, w, 0.0, w, 0.0, g, True,
                                       gwg, False, impl)
        dy2, dxall, ds2, _ = B.sim_bwd(y3, xall, None, None, None, sc, B.SCALE_MUL, 0.0,
                                       B.SIM_PLAIN, rank * n, r2, None, w, 0.0, w, 0.0, g, True,
                                       gwg, False, impl)
        dx, dy = dx1[0], dy2[0]
        if gwg:
            dx = dx + _reduce_scatter_sum(dxall[0], group)
            dy = dy + _reduce_scatter_sum(dyall[0], group)
        dls = None
        if ctx.needs_input_grad[2]:
            dls = B.scale_param_grad(ds1 + ds2, logit_scale, B.SCALE_MUL, 0.0)
        return (dx.to(in_dtype) if ctx.needs_input_grad[0] else None,
                dy.to(in_dtype) if ctx.needs_input_grad[1] else None, dls, None, None, None, None)


def sharded_local_clip_loss(image_local, text_local, logit_scale, group=None, backend=None,
                            gather_with_grad=False, impl=0):
    """open_clip `ClipLoss(local_loss=True, ...)` on this rank's shard (see _ShardedLocalCEFn)."""
    if backend is None:
        from . import ops as backend
    group = group if group is not None else dist.group.WORLD
    if not torch.is_tensor(logit_scale):
        logit_scale = torch.tensor(float(logit_scale), device=image_local.device)
    return _ShardedLocalCEFn.apply(image_local, text_local, logit_scale, group, backend,
                                   bool(gather_with_grad), int(impl))


class _ShardedModCEFn(torch.autograd.Function):
    """mean over the GLOBAL rows of LSE_j z_ij - z_ii with
    z = s [(F F^T) o (1 - S S^T) + diag(D)] (loss.py:229-242): rank r holds rows
    [r n, (r+1) n) of F, of the normalised semantic features S and of D; F and S are
    all-gathered, the local [n, N] block runs the MODULATED kernels with the positives at
    column offset r n, the partial dF of the gathered operand is reduce-scattered."""

    @staticmethod
    def forward(ctx, feats, sem_hat, paired, logit_scale, group, backend, grad_scale, impl):
        B = backend
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        in_dtype = feats.dtype
        f = feats.detach()
        if f.dtype not in (torch.float32, torch.bfloat16):
            f = f.float()
        f = f.contiguous()
        n = f.shape[0]
        N = n * world
        # both gathers go through NVLink peer memory when it is available (one store kernel +
        # one barrier each, as in _ShardedClipFn); NCCL all-gather otherwise
        u = sem_hat.detach().float().contiguous()
        pf = _peer_all_gather(B, f, "mod_f", group, barrier=False)
        pu = _peer_all_gather(B, u, "mod_u", group) if pf is not None else None
        if pf is not None and pu is not None:
            fall, uall = pf[0].unsqueeze(0), pu[0].unsqueeze(0)
            ctx.gens = ((pf[1], _SYMM_GEN[pf[1]]), (pu[1], _SYMM_GEN[pu[1]]))
        else:
            if pf is not None:   # the first gather was issued without its barrier: complete it
                dist.barrier(group=group)
            fall = _all_gather_cat(f, group).unsqueeze(0)
            uall = _all_gather_cat(u, group).unsqueeze(0)
            ctx.gens = None
        D = paired.detach().float().contiguous().reshape(1, n)
        sc = logit_scale.detach().reshape(1).float().contiguous()
        row_lse, _, _, diag = B.sim_fwd(f.unsqueeze(0), fall, u.unsqueeze(0), uall, D, sc,
                                        B.SCALE_MUL, 0.0, B.SIM_MODULATED, rank * n, False, impl)
        # (1/N) sum_i (r_i - z_ii) over the local rows, then one scalar all-reduce
        part = B.clip_loss_finalize(row_lse * 2.0, diag, None, None, None, N, sc, B.SCALE_MUL,
                                    0.0, 0.0)
        red = _all_reduce_scalar(B, part.reshape(1).clone(), "mod_loss", group)
        ctx.save_for_backward(f, fall, u, uall, D, sc, row_lse, logit_scale)
        ctx.cfg = (group, B, grad_scale, impl, in_dtype, rank, n, N)
        return red.reshape(())

    @staticmethod
    def backward(ctx, gout):
        f, fall, u, uall, D, sc, row_lse, logit_scale = ctx.saved_tensors
        group, B, grad_scale, impl, in_dtype, rank, n, N = ctx.cfg
        if ctx.gens is not None and any(_SYMM_GEN.get(k) != gen for k, gen in ctx.gens):
            # another forward reused the symmetric gather buffers since: gather again
            fall = _all_gather_cat(f, group).unsqueeze(0)
            uall = _all_gather_cat(u, group).unsqueeze(0)
        g = (gout.detach().float().reshape(1) * grad_scale).contiguous()
        nn_ = float(N)
        da, db, dscale, ddiag = B.sim_bwd(f.unsqueeze(0), fall, u.unsqueeze(0), uall, D, sc,
                                          B.SCALE_MUL, 0.0, B.SIM_MODULATED, rank * n, row_lse,
                                          None, 1.0 / nn_, 0.0, 1.0 / nn_, 0.0, g, True, True,
                                          True, impl)
        dfeats = dpaired = dls = None
        if ctx.needs_input_grad[0]:
            dfeats = (da[0] + _reduce_scatter_sum(db[0], group)).to(in_dtype)
        if ctx.needs_input_grad[2]:
            dpaired = ddiag[0]
        if ctx.needs_input_grad[3]:
            dist.all_reduce(dscale, group=group)
            dls = B.scale_param_grad(dscale, logit_scale, B.SCALE_MUL, 0.0)
        return dfeats, None, dpaired, dls, None, None, None, None


class ShardedAlignClipLoss(nn.Module):
    """open_clip-style ClipLoss (loss.py:67-132) of the GLOBAL batch from local shards: what
    `world_size > 1, gather_with_grad=True, local_loss=False` computes, without gathering the
    gradients through autograd and without materialising [N, N] logits."""

    def __init__(self, group=None, grad_scale: float = 1.0, backend=None):
        super().__init__()
        self.group, self.grad_scale, self.backend, self.impl = group, grad_scale, backend, 0

    def _ctx(self, ref):
        backend = self.backend
        if backend is None:
            from . import ops as backend
        group = self.group if self.group is not None else dist.group.WORLD
        return backend, group

    def forward(self, image_features, text_features, logit_scale, output_dict=False,
                semantic_features=None):
        backend, group = self._ctx(image_features)
        if not torch.is_tensor(logit_scale):
            logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
        loss = _sharded_ce_pair(image_features, text_features, logit_scale, group, backend,
                                self.grad_scale, self.impl)
        return {"contrastive_loss": loss} if output_dict else loss


class ShardedClipInModalityLoss(ShardedAlignClipLoss):
    """ClipInModalityLoss (loss.py:135-273; AlignCLIPLoss = alpha 1, beta 0.5, semantic) of the
    GLOBAL batch from local [n, d] feature shards and local [n, d_sem] semantic features (cfg5:
    N = 65536 over 8 GPUs).  The reference's own world_size > 1 branch never gathers the
    semantic features; this is the single-process formula on the concatenated
    batch, which is what the tests check it against."""

    def __init__(self, alpha=1.0, beta=0.5, nl_semantic_supervision=False, separate_text=True,
                 separate_image=False, group=None, grad_scale: float = 1.0, backend=None):
        super().__init__(group, grad_scale, backend)
        self.alpha, self.beta = alpha, beta
        self.nl_semantic_supervision = nl_semantic_supervision
        self.separate_text, self.separate_image = separate_text, separate_image

    def forward(self, image_features, text_features, logit_scale, output_dict=False,
                semantic_features=None):
        backend, group = self._ctx(image_features)
        if not torch.is_tensor(logit_scale):
            logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
        args = (group, backend, self.grad_scale, self.impl)
        if self.nl_semantic_supervision:
            if semantic_features is None:
                raise TypeError("nl_semantic_supervision=True needs semantic_features")
            sem = semantic_features.detach().float().unsqueeze(0)
            sem_hat, _, _ = backend.rownorm_fwd(sem, True, 0.0, backend.F32, False)
            paired = (image_features.float() * text_features.float()).sum(dim=-1)
            terms = []
            if self.separate_text:
                terms.append(_ShardedModCEFn.apply(text_features, sem_hat[0], paired,
                                                   logit_scale, *args))
            if self.separate_image:
                terms.append(_ShardedModCEFn.apply(image_features, sem_hat[0], paired,
                                                   logit_scale, *args))
            if not terms:
                raise UnboundLocalError("inModality_loss: neither separate_text nor "
                                        "separate_image is set (reference loss.py:249-255)")
            inModality_loss = self.beta * sum(terms[1:], terms[0])
        else:
            inModality_loss = self.beta * (
                _sharded_ce_pair(image_features, image_features, logit_scale, *args) +
                _sharded_ce_pair(text_features, text_features, logit_scale, *args))
        clip_loss = self.alpha * (2.0 * _sharded_ce_pair(image_features, text_features,
                                                         logit_scale, *args))
        total_loss = inModality_loss + clip_loss
        if output_dict:
            return {"total_loss": total_loss, "clip_loss": clip_loss,
                    "inModality_loss": inModality_loss}
        return total_loss


# ----------------------------------------------------------------------------------------------
# Retrieval over a gallery sharded across the ranks (last row)
# ----------------------------------------------------------------------------------------------
def _gather_ragged_counts(n_local: int, device, group):
    """[P] int64 shard sizes in rank order (shards may differ in length)."""
    world = dist.get_world_size(group)
    mine = torch.tensor([n_local], dtype=torch.int64, device=device)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    return torch.cat(parts)


def sharded_topk(scores_local: torch.Tensor, k: int, group=None, backend=None):
    """Exact top-k over a gallery whose columns are sharded across the ranks.

    scores_local [Q, n_r] fp32: the scores of the SAME Q queries (replicated) against this
    rank's gallery shard (rank r holds global columns [off_r, off_r + n_r), off_r = sum of the
    earlier shards).  Returns (indices int64 [Q, k] GLOBAL, values [Q, k]) sorted descending,
    identical on every rank and identical to `topk_rows` on the concatenated [Q, sum n_r] matrix
    (metrics/topk.py:43-48 on the full gallery): per-shard top-min(k, n_r), one all-gather of the
    [Q, k] candidates, then a k-way merge by the same exact kernel.  Ties: lowest global index
    first — candidates are laid out in rank order and each shard's list is already in
    (value desc, index asc) order, so candidate position order equals global-index order among
    equal values, which is the kernel's tie rule."""
    B = backend
    if B is None:
        from . import ops as B
    group = group if group is not None else dist.group.WORLD
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    s = scores_local.detach().float()
    if s.dim() == 1:
        s = s.unsqueeze(0)
    Q, n_r = s.shape
    counts = _gather_ragged_counts(n_r, s.device, group)
    total = int(counts.sum())
    if not 1 <= k <= total:
        raise RuntimeError("selected index k out of range")  # what torch.topk raises
    offset = int(counts[:rank].sum())
    kl = min(k, n_r)
    # fixed-size candidate slots so one all_gather serves ragged shards: unused slots hold
    # (-inf, sentinel index); a sentinel can only be selected when k exceeds the real entries,
    # which the range check above excludes (ties at -inf resolve to the lower position = real)
    cand_v = torch.full((Q, k), float("-inf"), dtype=torch.float32, device=s.device)
    cand_i = torch.full((Q, k), total, dtype=torch.int64, device=s.device)
    if kl > 0:
        idx, val = B.topk_rows(s.contiguous(), kl)
        cand_v[:, :kl] = val
        cand_i[:, :kl] = idx + offset
    all_v = _all_gather_cat(cand_v.unsqueeze(0), group)       # [P, Q, k]
    all_i = _all_gather_cat(cand_i.unsqueeze(0), group)
    # real entries of shard r first, its padding last: put every padding slot behind every real
    # one so that a real -inf score still wins the position tie-break against padding
    flat_v = all_v.permute(1, 0, 2).reshape(Q, world * k)
    flat_i = all_i.permute(1, 0, 2).reshape(Q, world * k)
    pad = flat_i == total
    order = torch.argsort(pad.to(torch.int8), dim=1, stable=True)
    flat_v = flat_v.gather(1, order).contiguous()
    flat_i = flat_i.gather(1, order)
    pos, val = B.topk_rows(flat_v, k)
    return flat_i.gather(1, pos), val


def sharded_top_k_score(text_local, owner_image_local, images_all, k=5, group=None,
                        backend=None):
    """top_k_score (metrics/topk.py:54-82) with the TEXT gallery sharded: every rank holds a
    slice `text_local` [n_r, d] of the stacked text embeddings with `owner_image_local` [n_r]
    (index of the image each text belongs to) and ALL image embeddings `images_all` [Q, d]
    (replicated).  An image is correct when one of its own texts is in its global top-k.
    Returns (accuracy, {"correct", "total"}) on every rank."""
    t = text_local.float()
    q = images_all.float()
    t = t / t.norm(dim=1, keepdim=True).clamp_min(1e-8)  # cosine_similarity eps (topk.py:44)
    q = q / q.norm(dim=1, keepdim=True).clamp_min(1e-8)
    group = group if group is not None else dist.group.WORLD
    idx, _ = sharded_topk(q @ t.T, k, group, backend)
    owners = _all_gather_ragged(owner_image_local.to(torch.int64), group)
    hit = (owners[idx] == torch.arange(q.shape[0], device=q.device).view(-1, 1)).any(dim=1)
    correct = int(hit.sum().item())
    total = q.shape[0]
    return (correct / total if total > 0 else 0.0), {"correct": correct, "total": total}


def _all_gather_ragged(t: torch.Tensor, group):
    """concatenate 1-D shards of different lengths in rank order."""
    counts = _gather_ragged_counts(t.shape[0], t.device, group)
    m = int(counts.max())
    padded = torch.zeros((1, m), dtype=t.dtype, device=t.device)
    padded[0, :t.shape[0]] = t
    allp = _all_gather_cat(padded, group)                     # [P, m]
    return torch.cat([allp[r, :int(counts[r])] for r in range(allp.shape[0])])


def sharded_paired_ranks(image_local, text_local, logit_scale=None, group=None, backend=None):
    """get_clip_metrics' ranks (AlignCLIP/main/train.py:338-356) on a row-sharded batch: rank r
    holds rows [r n, (r+1) n) of both modalities.  The text shard is all-gathered, the local
    [n, N] block of `logit_scale * image @ text.T` feeds `mm25_rank_both` once: the row counts
    (image -> text) are complete locally, the column counts (text -> image) are partial sums
    over the local rows and are all-reduced (integers: exact, order-free).  The column targets
    (the diagonal) are all-gathered first.  Returns (i2t_local [n], t2i_local [n]) int32: the
    ranks of the LOCAL pairs."""
    B = backend
    if B is None:
        from . import ops as B
    group = group if group is not None else dist.group.WORLD
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    x, y = image_local.detach(), text_local.detach()
    n = x.shape[0]
    N = n * world
    yall = _all_gather_cat(y.contiguous(), group)
    s = (logit_scale * x @ yall.t()) if logit_scale is not None else x @ yall.t()
    s = s.float().contiguous()
    cols = torch.arange(rank * n, (rank + 1) * n, device=s.device, dtype=torch.int64)
    diag_local = s.gather(1, cols.view(-1, 1)).view(-1).contiguous()
    diag_all = _all_gather_cat(diag_local, group)
    rows_all = torch.arange(N, device=s.device, dtype=torch.int64)
    i2t = torch.zeros((n,), dtype=torch.int32, device=s.device)
    t2i = torch.zeros((N,), dtype=torch.int32, device=s.device)
    B.rank_both(s, rank * n, diag_local, cols, diag_all, rows_all, i2t, t2i)
    dist.all_reduce(t2i, group=group)
    return i2t, t2i[rank * n:(rank + 1) * n].clone()


# ----------------------------------------------------------------------------------------------
# CLOOB (reference losses/cloobLoss.py) on the row-sharded global batch
# ----------------------------------------------------------------------------------------------
class _AllGatherRowsFn(torch.autograd.Function):
    """[n, d] per rank -> [P n, d]; backward reduce-scatters (sums) the gradient of the gathered
    tensor so every rank receives the gradient of ITS rows from all ranks' partial losses — what
    `torch.distributed.nn.all_gather` + cat does in loss.py:49-51, in one collective."""

    @staticmethod
    def forward(ctx, t, group):
        ctx.group = group
        return _all_gather_cat(t.detach(), group)

    @staticmethod
    def backward(ctx, g):
        return _reduce_scatter_sum(g.detach().contiguous(), ctx.group), None


class _AllReduceSumFn(torch.autograd.Function):
    """sum of a scalar part over the ranks; d(sum)/d(part) = 1 on every rank."""

    @staticmethod
    def forward(ctx, t, group):
        out = t.detach().clone()
        dist.all_reduce(out, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        return g, None


def sharded_cloob_loss(image_local, text_local, inv_tau, scale_hopfield, group=None,
                       backend=None, grad_scale: float = 1.0):
    """cloob_loss (cloobLoss.py:10-21) of the GLOBAL batch from the local [n, d] shards.

    The Hopfield softmax runs over the STORED set, so both modalities are gathered (as
    cloob-training/train.py:241-242 does); each rank then retrieves for its own n state
    patterns against all N stored ones (rectangular [n, N] retrievals, complete per row), the
    retrieved text-side patterns are gathered once more as the columns of InfoLOOB, and the
    rank's rows contribute `sum_local (LSE'_i - k_ii) / N` with the positives at column offset
    r n.  The scalar parts are summed over the ranks.  Every gather is differentiable
    (reduce-scatter of the gradient), so the gradients are those of the global loss w.r.t. the
    local shards, times `grad_scale` (cf. ShardedClipLoss)."""
    from .losses.cloobLoss import _HopfieldFn, _InfoLoobFn
    group = group if group is not None else dist.group.WORLD
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = image_local.shape[0]
    N, off = n * world, rank * n
    beta = float(scale_hopfield)
    x_all = _AllGatherRowsFn.apply(image_local, group)
    y_all = _AllGatherRowsFn.apply(text_local, group)
    # rows = local state patterns, softmax over all stored patterns (cloobLoss.py:35-40)
    u_xx = _HopfieldFn.apply(image_local, x_all, beta, backend)   # state image, stored image
    u_yy = _HopfieldFn.apply(text_local, y_all, beta, backend)    # state text,  stored text
    u_xy = _HopfieldFn.apply(text_local, x_all, beta, backend)    # state text,  stored image
    u_yx = _HopfieldFn.apply(image_local, y_all, beta, backend)   # state image, stored text
    u_xy_all = _AllGatherRowsFn.apply(u_xy, group)
    u_yx_all = _AllGatherRowsFn.apply(u_yx, group)
    part_img = _InfoLoobFn.apply(u_xx, u_xy_all, float(inv_tau), off, N, backend)  # :18
    part_txt = _InfoLoobFn.apply(u_yy, u_yx_all, float(inv_tau), off, N, backend)  # :19
    part = (part_img + part_txt) / 2                                               # :20
    if grad_scale != 1.0:  # forward value unchanged, gradients scaled
        part = part * grad_scale + (part * (1.0 - grad_scale)).detach()
    return _AllReduceSumFn.apply(part, group)


class ShardedCLOOBLoss(nn.Module):
    """CLOOBLoss(inv_tau, scale_hopfield, device) of the GLOBAL batch from local shards."""

    def __init__(self, inv_tau, scale_hopfield, device=None, group=None, grad_scale: float = 1.0,
                 backend=None):
        super().__init__()
        self.inv_tau, self.scale_hopfield, self.device = inv_tau, scale_hopfield, device
        self.group, self.grad_scale, self.backend = group, grad_scale, backend

    def forward(self, image_features, text_features):
        return sharded_cloob_loss(image_features, text_features, self.inv_tau,
                                  self.scale_hopfield, self.group, self.backend, self.grad_scale)


# ----------------------------------------------------------------------------------------------
# vCLIP on the row-sharded batch: the sample-batched CLIP term (clipLoss.py:24-79) shards like
# cfg2 per sample slice; the samplers are row-parallel (replicas only, no collective)
# ----------------------------------------------------------------------------------------------
def _gather_dim1(t: torch.Tensor, group) -> torch.Tensor:
    """[S, n, d] per rank -> [S, P n, d] (rows in rank order)."""
    return _all_gather_cat(t.transpose(0, 1).contiguous(), group).transpose(0, 1).contiguous()


def _reduce_scatter_dim1(t: torch.Tensor, group) -> torch.Tensor:
    """[S, P n, d] partial per rank -> sum over ranks of the local [S, n, d] rows."""
    return _reduce_scatter_sum(t.transpose(0, 1).contiguous(), group).transpose(0, 1).contiguous()


class _ShardedClipSampledFn(torch.autograd.Function):
    """loss [S] of S independent GLOBAL CLIP problems from local [S, n, d] sample shards: the
    algorithm of _ShardedClipFn with the batch dimension of the fused kernels = S (one gather,
    one statistics exchange, one reduce-scatter for all slices)."""

    @staticmethod
    def forward(ctx, img, txt, logits_scale, eps_ls, clamp_max, group, backend, grad_scale, impl):
        B = backend
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        in_dtype = img.dtype
        if in_dtype not in (torch.float32, torch.bfloat16):
            img, txt = img.float(), txt.float()
        if txt.dtype != img.dtype:
            txt = txt.to(img.dtype)
        cdt = B.BF16 if img.dtype == torch.bfloat16 else B.F32
        S, n, d = img.shape
        N = n * world
        want_cs = eps_ls != 0.0
        xhat, invx, sx = B.rownorm_fwd(img.contiguous(), True, _EPS_NORM, cdt, want_cs)
        yhat, invy, sy = B.rownorm_fwd(txt.contiguous(), True, _EPS_NORM, cdt, want_cs)
        scale32 = logits_scale.detach().reshape(1).to(torch.float32).contiguous()
        yall = _gather_dim1(yhat, group)
        row_lse, cmax, csum, diag = B.sim_fwd(xhat, yall, None, None, None, scale32,
                                              B.SCALE_LOG, clamp_max, B.SIM_PLAIN, rank * n,
                                              True, impl, 1.01)
        extra = [sx.reshape(S, d), sy.reshape(S, d)] if want_cs else []
        payload = torch.cat([cmax.reshape(S, N), csum.reshape(S, N)] + extra, dim=1)
        gathered = _all_gather_cat(payload.unsqueeze(0), group)          # [P, S, 2N (+ 2d)]
        col_lse = B.col_merge(gathered[:, :, :N].reshape(world, S * N).contiguous(),
                              gathered[:, :, N:2 * N].reshape(world, S * N).contiguous()
                              ).reshape(S, N)
        part = B.clip_loss_finalize(row_lse, diag,
                                    col_lse[:, rank * n:(rank + 1) * n].contiguous(), None, None,
                                    N, scale32, B.SCALE_LOG, clamp_max, eps_ls)
        red = part.reshape(S).clone()
        dist.all_reduce(red, group=group)
        if want_cs:
            sxg = gathered[:, :, 2 * N:2 * N + d].sum(dim=0)
            syg = gathered[:, :, 2 * N + d:2 * N + 2 * d].sum(dim=0)
            s = scale32.clamp(max=clamp_max).exp().reshape(())
            red = red - (eps_ls / float(N) ** 2) * s * (sxg * syg).sum(dim=1)
        ctx.save_for_backward(xhat, yhat, yall, invx, invy, row_lse, col_lse, scale32,
                              logits_scale)
        ctx.cfg = (eps_ls, clamp_max, group, B, grad_scale, impl, in_dtype, rank, n, N)
        return red

    @staticmethod
    def backward(ctx, gout):
        (xhat, yhat, yall, invx, invy, row_lse, col_lse, scale32,
         logits_scale) = ctx.saved_tensors
        eps_ls, clamp_max, group, B, grad_scale, impl, in_dtype, rank, n, N = ctx.cfg
        nn_ = float(N)
        g = (gout.detach().to(torch.float32).reshape(-1) * grad_scale).contiguous()
        da, db, dscale, _ = B.sim_bwd(xhat, yall, None, None, None, scale32, B.SCALE_LOG,
                                      clamp_max, B.SIM_PLAIN, rank * n, row_lse, col_lse,
                                      0.5 / nn_, 0.5 / nn_, (1.0 - eps_ls) / nn_,
                                      eps_ls / (nn_ * nn_), g, True, True, False, impl, 1.01)
        dy_local = _reduce_scatter_dim1(db, group)
        out_dt = B.BF16 if in_dtype == torch.bfloat16 else B.F32
        dimg = B.rownorm_bwd(da, xhat, invx, True, out_dt, None).to(in_dtype)
        dtxt = B.rownorm_bwd(dy_local, yhat, invy, True, out_dt, None).to(in_dtype)
        dls = None
        if ctx.needs_input_grad[2]:
            dscale = dscale.reshape(1).clone()
            dist.all_reduce(dscale, group=group)
            dls = B.scale_param_grad(dscale, logits_scale, B.SCALE_LOG, clamp_max)
        return dimg, dtxt, dls, None, None, None, None, None, None


def make_sharded_vclip_loss(group=None, backend=None, grad_scale: float = 1.0, **vclip_kwargs):
    """VClipLoss (losses/vclipLoss.py:43-162) of the GLOBAL batch from local distribution
    shards: sampling stays local (row-parallel), the CLIP term runs on the sharded batch
    (`ShardedClipLoss.forward` / `.forward_sampled`), and the KL / variance-regulariser means
    over the batch become global means (differentiable all-reduce / P).  Returns the module."""
    from .losses.vclipLoss import VClipLoss

    class ShardedVClipLoss(VClipLoss):
        def __init__(self):
            super().__init__(**vclip_kwargs)
            self.clip_loss = ShardedClipLoss(self.clip_loss.label_smoothing, group, grad_scale,
                                             backend)
            self._group = group

        def forward(self, image_distribution, text_distribution, image_vars, text_vars,
                    logits_scale, is_train, kl_weight_override=None):
            out = super().forward(image_distribution, text_distribution, image_vars, text_vars,
                                  logits_scale, is_train, kl_weight_override)
            grp = self._group if self._group is not None else dist.group.WORLD
            world = dist.get_world_size(grp)
            for key in ("image_kl_loss", "text_kl_loss", "var_reg"):  # local mean -> global mean
                out[key] = _AllReduceSumFn.apply(out[key], grp) / world
            kl_weight = kl_weight_override if kl_weight_override is not None else self.kl_weight
            out["total_loss"] = (self.clip_weight * out["clip_loss"] + 0.5 * kl_weight * (
                out["image_kl_loss"] + out["text_kl_loss"]) + self.var_reg_weight * out["var_reg"])
            return out

    return ShardedVClipLoss()

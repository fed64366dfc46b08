"""TEST INFRASTRUCTURE: a dense torch-CPU stand-in for `multimodal_2025_b200.ops`, used only to
exercise the collective / merge logic of dist.ShardedClipLoss under gloo (no GPU in CI).
It restates each op's contract (include/mm25.h) with materialised logits."""
import torch

F32, BF16 = 0, 1
SIM_PLAIN, SIM_MASKDIAG, SIM_MODULATED = 0, 1, 2
SCALE_MUL, SCALE_LOG = 0, 1


def _s(scale, kind, clamp):
    return scale.clamp(max=clamp).exp() if kind == SCALE_LOG else scale


def rownorm_fwd(x, normalize, eps, out_dtype, want_colsum):
    xf = x.float()
    inv = 1.0 / xf.norm(dim=-1).clamp_min(eps) if normalize else torch.ones(x.shape[:2])
    xhat = (xf * inv.unsqueeze(-1)).to(torch.float32 if out_dtype == F32 else torch.bfloat16)
    cs = xhat.float().sum(dim=1) if want_colsum else torch.empty(0)
    return xhat, inv, cs


def sim_fwd(a, b, u, v, diag_add, scale, scale_kind, clamp_max, mode, diag_offset, want_cols,
            impl, dot_bound=0.0):
    M = a.shape[1]
    idx = torch.arange(M)
    base = a.float() @ b.float().transpose(1, 2)
    if mode == SIM_MODULATED:
        base = base * (1.0 - u.float() @ v.float().transpose(1, 2))
        base[:, idx, idx + diag_offset] += diag_add
    z = _s(scale, scale_kind, clamp_max) * base
    diag = z[:, idx, idx + diag_offset].clone()
    if mode == SIM_MASKDIAG:  # InfoLOOB: the positive is replaced by -10000 inside the LSE
        z = z.clone()
        z[:, idx, idx + diag_offset] = -10000.0
    cmax = z.max(dim=1).values
    csum = (z - cmax.unsqueeze(1)).exp().sum(dim=1)
    return torch.logsumexp(z, dim=2), cmax, csum, diag


def col_merge(col_max, col_sum):
    m = col_max.max(dim=0).values
    return m + (col_sum * (col_max - m).exp()).sum(dim=0).log()


def clip_loss_finalize(row_lse, diag, col_lse, sx, sy, n_norm, scale, scale_kind, clamp_max,
                       eps_ls):
    n = float(n_norm)
    out = row_lse.double().sum(1) / (2 * n) - (1 - eps_ls) / n * diag.double().sum(1)
    if col_lse is not None:
        out = out + col_lse.double().sum(1) / (2 * n)
    if eps_ls != 0 and sx is not None:
        out = out - eps_ls / n ** 2 * _s(scale, scale_kind, clamp_max).double() * \
            (sx.double() * sy.double()).sum(1)
    return out.float()


def sim_bwd(a, b, u, v, diag_add, scale, scale_kind, clamp_max, mode, diag_offset, row_lse,
            col_lse, w_row, w_col, w_diag, w_all, gout, need_da, need_db, need_ddiag, impl,
            dot_bound=0.0):
    s = _s(scale, scale_kind, clamp_max)
    M = a.shape[1]
    idx = torch.arange(M)
    dot = a.float() @ b.float().transpose(1, 2)
    chain = 1.0
    if mode == SIM_MODULATED:
        chain = 1.0 - u.float() @ v.float().transpose(1, 2)
        dot = dot * chain
        dot[:, idx, idx + diag_offset] += diag_add
    z = s * dot
    if mode == SIM_MASKDIAG:
        z = z.clone()
        z[:, idx, idx + diag_offset] = -10000.0
    G = w_row * (z - row_lse.unsqueeze(2)).exp() - w_all
    if col_lse is not None:
        G = G + w_col * (z - col_lse.unsqueeze(1)).exp()
    G[:, idx, idx + diag_offset] -= w_diag
    go = gout.reshape(-1, 1, 1) if gout is not None else 1.0
    G = G * go
    da = s * ((G * chain) @ b.float())
    db = s * ((G * chain).transpose(1, 2) @ a.float())
    dscale = (G * dot).sum().reshape(1)
    ddiag = s * G[:, idx, idx + diag_offset] if need_ddiag else torch.empty(0)
    return da, db, dscale, ddiag


def rownorm_bwd(g, xhat, inv_norm, normalize, out_dtype, gscale):
    h = xhat.float()
    if normalize:
        dx = (g - h * (h * g).sum(-1, keepdim=True)) * inv_norm.unsqueeze(-1)
    else:
        dx = g
    return dx.to(torch.float32 if out_dtype == F32 else torch.bfloat16)


def scale_param_grad(dscale, scale, scale_kind, clamp_max):
    if scale_kind == SCALE_MUL:
        return dscale.reshape(scale.shape)
    l = scale.detach().float()
    return (dscale.reshape(()) * torch.where(l <= clamp_max, l.exp(), torch.zeros_like(l))
            ).reshape(scale.shape)


def _order_key(v):
    """int64 keys whose order is torch's sort order of fp32 (NaN largest, -0.0 < +0.0 as the
    kernels' bit-pattern keys have it)."""
    u = v.contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    key = torch.where(u >= 0x80000000, 0xFFFFFFFF - u, u | 0x80000000)
    return torch.where(torch.isnan(v), torch.full_like(key, 0xFFFFFFFF), key)


def rank_both(scores, row0, row_tval, row_tcol, col_tval, col_trow, row_rank, col_rank):
    """include/mm25.h mm25_rank_both, dense: counts of entries ahead of the row / column target
    in the stable descending order, ADDED into the int32 outputs."""
    Q, N = scores.shape
    key = _order_key(scores.float())
    if row_rank is not None:
        tk = _order_key(row_tval).view(-1, 1)
        pos = torch.arange(N).view(1, -1)
        ahead = (key > tk) | ((key == tk) & (pos < row_tcol.view(-1, 1)))
        row_rank += ahead.sum(dim=1).to(torch.int32)
    if col_rank is not None:
        tk = _order_key(col_tval).view(1, -1)
        pos = (row0 + torch.arange(Q)).view(-1, 1)
        ahead = (key > tk) | ((key == tk) & (pos < col_trow.view(1, -1)))
        col_rank += ahead.sum(dim=0).to(torch.int32)


def topk_rows(scores, k):
    """include/mm25.h mm25_topk_rows, dense: stable descending sort, first k."""
    s = scores.float()
    if s.dim() == 1:
        s = s.unsqueeze(0)
    key = _order_key(s)
    idx = torch.argsort(key, dim=1, descending=True, stable=True)[:, :k]
    return idx, s.gather(1, idx)


IMPL_AUTO, IMPL_SIMT, IMPL_TC = 0, 1, 2


def rowdot(x, y):
    return (x * y).sum(dim=-1)


def hopfield_bwd_pass(a, b, a2, b2, bm1, bm2, beta, lse, D, stat_by_col):
    """include/mm25.h mm25_hopfield_bwd_pass, dense."""
    S = a @ b.T
    dP = a2 @ b2.T
    l = lse.reshape(1, -1) if stat_by_col else lse.reshape(-1, 1)
    Dq = D.reshape(1, -1) if stat_by_col else D.reshape(-1, 1)
    p = (beta * S - l).exp()
    out = (beta * p * (dP - Dq)) @ bm1
    if bm2 is not None:
        out = out + p @ bm2
    return out

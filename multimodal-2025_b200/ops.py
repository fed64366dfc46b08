"""torch.library custom ops (`mm25::*`) over the C-ABI of libmm25.so.

CUDA dispatch only: CPU tensors raise (there is no fallback).  Every op allocates its outputs
and workspace through PyTorch's caching allocator and launches on the current stream, so the
kernels never allocate (include/mm25.h conventions).
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
from torch import Tensor

from . import _lib

F32, BF16 = 0, 1
SIM_PLAIN, SIM_MASKDIAG, SIM_MODULATED = 0, 1, 2
SCALE_MUL, SCALE_LOG = 0, 1
IMPL_AUTO, IMPL_SIMT, IMPL_TC = 0, 1, 2
# |xhat_i . yhat_j| for rows normalised by rownorm_fwd (bf16 rounding included)
DOT_BOUND_NORMALIZED = 1.01


def _dt(t: Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError("mm25 ops take float32 or bfloat16 embeddings, got %s" % t.dtype)


def _tdt(code: int) -> torch.dtype:
    return torch.float32 if code == F32 else torch.bfloat16


def _p(t: Optional[Tensor]):
    return None if t is None else t.data_ptr()


def _stream(t: Tensor):
    return torch.cuda.current_stream(t.device).cuda_stream


def _cuda3(t: Tensor, name: str) -> Tensor:
    if not t.is_cuda:
        raise _lib.MM25Error("%s must be a CUDA tensor: this path has no CPU fallback" % name)
    if t.dim() != 3:
        raise ValueError("%s must be [B, M, d]" % name)
    return t.contiguous()


def _f32c(t: Optional[Tensor]) -> Optional[Tensor]:
    if t is None:
        return None
    return t.detach().to(torch.float32).contiguous()


@torch.library.custom_op("mm25::rownorm_fwd", mutates_args=(), device_types="cuda")
def rownorm_fwd(x: Tensor, normalize: bool, eps: float, out_dtype: int,
                want_colsum: bool) -> Tuple[Tensor, Tensor, Tensor]:
    """F.normalize(x, dim=-1) (+ inverse norms, + column sums). Ref: clipLoss.py:28-29,95-96."""
    x = _cuda3(x, "x")
    B, M, d = x.shape
    with torch.cuda.device(x.device):
        xhat = torch.empty((B, M, d), dtype=_tdt(out_dtype), device=x.device)
        inv = torch.empty((B, M), dtype=torch.float32, device=x.device)
        ncs = int(_lib.lib().mm25_rownorm_colsum_floats(B, M, d)) if want_colsum else 0
        cs = torch.empty((ncs,), dtype=torch.float32, device=x.device)
        rc = _lib.lib().mm25_rownorm_fwd(_p(x), _dt(x), _p(xhat), out_dtype, _p(inv),
                                         _p(cs) if want_colsum else None, B, M, d,
                                         int(normalize), eps, _stream(x))
    _lib.check(rc, "mm25_rownorm_fwd")
    return xhat, inv, (cs[:B * d].reshape(B, d).clone() if want_colsum else cs)


@torch.library.custom_op("mm25::rownorm_bwd", mutates_args=(), device_types="cuda")
def rownorm_bwd(g: Tensor, xhat: Tensor, inv_norm: Tensor, normalize: bool, out_dtype: int,
                gscale: Optional[Tensor]) -> Tensor:
    g = _cuda3(g, "g")
    B, M, d = g.shape
    xhat = xhat.contiguous()
    with torch.cuda.device(g.device):
        dx = torch.empty((B, M, d), dtype=_tdt(out_dtype), device=g.device)
        rc = _lib.lib().mm25_rownorm_bwd(_p(g), _p(xhat), _dt(xhat), _p(inv_norm), _p(dx),
                                         out_dtype, B, M, d, int(normalize), _p(gscale), 0,
                                         _stream(g))
    _lib.check(rc, "mm25_rownorm_bwd")
    return dx


def _sim_fwd_impl(a, b, u, v, diag_add, scale, scale_kind, clamp_max, mode, diag_offset,
                  want_cols, impl, dot_bound, keep):
    a = _cuda3(a, "a")
    b = _cuda3(b, "b")
    if a.dtype != b.dtype:
        raise TypeError("a and b must share a dtype")
    B, M, d = a.shape
    N = b.shape[1]
    d2 = 0 if u is None else u.shape[-1]
    dev = a.device
    L = _lib.lib()
    with torch.cuda.device(dev):
        row_lse = torch.empty((B, M), dtype=torch.float32, device=dev)
        diag = torch.zeros((B, M), dtype=torch.float32, device=dev)
        cshape = (B, N) if want_cols else (0,)
        col_max = torch.empty(cshape, dtype=torch.float32, device=dev)
        col_sum = torch.empty(cshape, dtype=torch.float32, device=dev)
        wsb = L.mm25_sim_fwd_workspace_bytes(B, M, N, d, _dt(a), impl)
        if mode == SIM_MODULATED and impl != IMPL_SIMT and not want_cols:
            wsb = max(wsb, L.mm25_sim_mod_workspace_bytes(B, M, N, d, d2, _dt(a), 0))
        ws = torch.empty((max(int(wsb), 16),), dtype=torch.uint8, device=dev)
        e_keep = torch.empty((B * M, N) if keep else (0,), dtype=torch.bfloat16, device=dev)
        e_flag = torch.zeros((1,) if keep else (0,), dtype=torch.int32, device=dev)
        rc = L.mm25_sim_fwd(_p(a), _p(b), _dt(a), _p(u), _p(v), _p(diag_add), B, M, N, d, d2,
                            _p(scale), scale_kind, clamp_max, mode, diag_offset, _p(row_lse),
                            _p(col_max) if want_cols else None,
                            _p(col_sum) if want_cols else None, _p(diag), _p(ws), ws.numel(),
                            impl, dot_bound, _p(e_keep) if keep else None,
                            _p(e_flag) if keep else None, _stream(a))
    _lib.check(rc, "mm25_sim_fwd")
    return row_lse, col_max, col_sum, diag, e_keep, e_flag


@torch.library.custom_op("mm25::sim_fwd", mutates_args=(), device_types="cuda")
def sim_fwd(a: Tensor, b: Tensor, u: Optional[Tensor], v: Optional[Tensor],
            diag_add: Optional[Tensor], scale: Tensor, scale_kind: int, clamp_max: float,
            mode: int, diag_offset: int, want_cols: bool,
            impl: int, dot_bound: float = 0.0) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Fused similarity + online row/column LSE; returns (row_lse, col_max, col_sum, diag).
    `dot_bound`: bound on |a_i.b_j| if the caller knows one (include/mm25.h), else 0."""
    return _sim_fwd_impl(a, b, u, v, diag_add, scale, scale_kind, clamp_max, mode, diag_offset,
                         want_cols, impl, dot_bound, False)[:4]


@torch.library.custom_op("mm25::sim_fwd_keep", mutates_args=(), device_types="cuda")
def sim_fwd_keep(a: Tensor, b: Tensor, scale: Tensor, scale_kind: int, clamp_max: float,
                 diag_offset: int, want_cols: bool, impl: int, dot_bound: float
                 ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    """PLAIN-mode sim_fwd that also keeps E = exp(z - shift) for the backward (include/mm25.h
    `e_keep`): returns (row_lse, col_max, col_sum, diag, e_keep [B*M, N] bf16, e_flag [1] int32).
    Pass the last two to sim_bwd / sim_bwd_rs, which consume e_keep."""
    return _sim_fwd_impl(a, b, None, None, None, scale, scale_kind, clamp_max, SIM_PLAIN,
                         diag_offset, want_cols, impl, dot_bound, True)


def keep_e_bytes_limit(M: int = 0, N: int = 0, d: int = 0) -> int:
    """Largest E buffer (bytes) the losses keep for the backward.  `MM25_KEEP_E_GB` overrides.

    Default: 6 GiB for shapes the fused-transform GEMM covers (csrc/gemm_te.cu: M, N, d multiples
    of 256) — the backward then reads E inside the GEMM's operand load and neither recomputes the
    S tiles nor makes a separate E -> G' pass; 0 (always recompute) for other shapes, where the
    kept-E path measured no better than the recompute (profiles/r1_keep_e.md)."""
    import os
    env = os.environ.get("MM25_KEEP_E_GB")
    if env is not None:
        return int(float(env) * (1 << 30))
    if M > 0 and M % 256 == 0 and N % 256 == 0 and d % 256 == 0 and \
            os.environ.get("MM25_TE", "1") != "0":
        return 6 << 30
    return 0


# e_keep is scratch owned by the loss's autograd node and consumed here (never an autograd-visible
# tensor), so it is not declared in mutates_args (torch cannot version-bump a defaulted argument).
@torch.library.custom_op("mm25::sim_bwd", mutates_args=(), device_types="cuda")
def sim_bwd(a: Tensor, b: Tensor, u: Optional[Tensor], v: Optional[Tensor],
            diag_add: Optional[Tensor], scale: Tensor, scale_kind: int, clamp_max: float,
            mode: int, diag_offset: int, row_lse: Tensor, col_lse: Optional[Tensor],
            w_row: float, w_col: float, w_diag: float, w_all: float, gout: Optional[Tensor],
            need_da: bool, need_db: bool, need_ddiag: bool,
            impl: int, dot_bound: float = 0.0, e_keep: Optional[Tensor] = None,
            e_flag: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Fused backward with tile recompute; returns (da, db, dscale[1], ddiag), all fp32.
    `e_keep`, `e_flag`: the pair sim_fwd_keep returned (consumed), or None to recompute.

    dscale is dL/d(multiplicative scale) = sum_ij G_ij z_ij / s (already times gout)."""
    a = _cuda3(a, "a")
    b = _cuda3(b, "b")
    B, M, d = a.shape
    N = b.shape[1]
    d2 = 0 if u is None else u.shape[-1]
    dev = a.device
    L = _lib.lib()
    with torch.cuda.device(dev):
        da = torch.empty((B, M, d) if need_da else (0,), dtype=torch.float32, device=dev)
        db = torch.empty((B, N, d) if need_db else (0,), dtype=torch.float32, device=dev)
        ddiag = torch.zeros((B, M) if need_ddiag else (0,), dtype=torch.float32, device=dev)
        nparts = int(L.mm25_sim_bwd_nparts(B, M, N, d, _dt(a), impl))
        with_e = e_keep is not None and e_keep.numel() > 0 and e_flag is not None
        wsb = L.mm25_sim_bwd_workspace_bytes(B, M, N, d, _dt(a), impl, int(with_e))
        if mode == SIM_MODULATED and impl != IMPL_SIMT and w_col == 0.0:
            wsb = max(wsb, L.mm25_sim_mod_workspace_bytes(B, M, N, d, d2, _dt(a), 1))
            nparts = max(nparts, int(L.mm25_sim_mod_bwd_nparts(B, M, N, d, d2, _dt(a))))
        parts = torch.zeros((max(nparts, 1),), dtype=torch.float32, device=dev)
        dscale = torch.empty((1,), dtype=torch.float32, device=dev)
        ws = torch.empty((max(int(wsb), 16),), dtype=torch.uint8, device=dev)
        rc = L.mm25_sim_bwd(_p(a), _p(b), _dt(a), _p(u), _p(v), _p(diag_add), B, M, N, d, d2,
                            _p(scale), scale_kind, clamp_max, mode, diag_offset, _p(row_lse),
                            _p(col_lse), w_row, w_col, w_diag, w_all, _p(gout),
                            _p(da) if need_da else None, _p(db) if need_db else None, _p(parts),
                            _p(ddiag) if need_ddiag else None, _p(ws), ws.numel(), impl,
                            dot_bound, _p(e_keep) if with_e else None,
                            _p(e_flag) if with_e else None, _stream(a))
        _lib.check(rc, "mm25_sim_bwd")
        rc = L.mm25_sum_parts(_p(parts), parts.numel(), 1.0, _p(dscale), _stream(a))
    _lib.check(rc, "mm25_sum_parts")
    return da, db, dscale, ddiag


def sim_split_eligible(B: int, M: int, N: int, d: int) -> bool:
    """True when the fp32-grade tensor-core path covers the shape (include/mm25.h)."""
    return bool(_lib.lib().mm25_sim_split_eligible(B, M, N, d))


@torch.library.custom_op("mm25::split_bf16x2", mutates_args=(), device_types="cuda")
def split_bf16x2(x: Tensor, role: int) -> Tuple[Tensor, Tensor, Tensor]:
    """fp32 [B, rows, d] -> (out4 [B, rows, W d], hi, lo [B, rows, d]) bf16: the two-term split
    x = hi + lo laid out for the K = W d sweep (W = mm25_split_width() = 3: role 0 = row operand
    [hi hi lo], role 1 = column operand [hi lo hi]; include/mm25.h mm25_split_bf16x2)."""
    x = _cuda3(x, "x")
    if x.dtype != torch.float32:
        raise TypeError("split_bf16x2 takes fp32 rows")
    B, R, d = x.shape
    with torch.cuda.device(x.device):
        W = int(_lib.lib().mm25_split_width())
        out4 = torch.empty((B, R, W * d), dtype=torch.bfloat16, device=x.device)
        hi = torch.empty((B, R, d), dtype=torch.bfloat16, device=x.device)
        lo = torch.empty((B, R, d), dtype=torch.bfloat16, device=x.device)
        rc = _lib.lib().mm25_split_bf16x2(_p(x), B * R, d, role, _p(out4), _p(hi), _p(lo),
                                          _stream(x))
    _lib.check(rc, "mm25_split_bf16x2")
    return out4, hi, lo


@torch.library.custom_op("mm25::sim_bwd_split", mutates_args=(), device_types="cuda")
def sim_bwd_split(a4: Tensor, b4: Tensor, a_hi: Tensor, a_lo: Tensor, b_hi: Tensor, b_lo: Tensor,
                  scale: Tensor, scale_kind: int, clamp_max: float, diag_offset: int,
                  row_lse: Tensor, col_lse: Optional[Tensor], w_row: float, w_col: float,
                  w_diag: float, w_all: float, gout: Optional[Tensor], need_da: bool,
                  need_db: bool, dot_bound: float,
                  mode: int = 0) -> Tuple[Tensor, Tensor, Tensor]:
    """fp32-grade PLAIN / MASKDIAG backward on the split operands (include/mm25.h mm25_sim_bwd_split);
    returns (da [B, M, d], db [B, N, d], dscale[1]) fp32."""
    B, M, d4 = a4.shape
    N = b4.shape[1]
    d = d4 // int(_lib.lib().mm25_split_width())
    dev = a4.device
    L = _lib.lib()
    with torch.cuda.device(dev):
        da = torch.empty((B, M, d) if need_da else (0,), dtype=torch.float32, device=dev)
        db = torch.empty((B, N, d) if need_db else (0,), dtype=torch.float32, device=dev)
        parts = torch.zeros((max(int(L.mm25_sim_bwd_split_nparts(B, M, N)), 1),),
                            dtype=torch.float32, device=dev)
        dscale = torch.empty((1,), dtype=torch.float32, device=dev)
        ws = torch.empty((max(int(L.mm25_sim_bwd_split_workspace_bytes(M, N)), 16),),
                         dtype=torch.uint8, device=dev)
        rc = L.mm25_sim_bwd_split(_p(a4), _p(b4), _p(a_hi), _p(a_lo), _p(b_hi), _p(b_lo), B, M, N,
                                  d, _p(scale), scale_kind, clamp_max, diag_offset, _p(row_lse),
                                  _p(col_lse), w_row, w_col, w_diag, w_all, _p(gout),
                                  _p(da) if need_da else None, _p(db) if need_db else None,
                                  _p(parts), _p(ws), ws.numel(), dot_bound, mode, _stream(a4))
        _lib.check(rc, "mm25_sim_bwd_split")
        rc = L.mm25_sum_parts(_p(parts), parts.numel(), 1.0, _p(dscale), _stream(a4))
    _lib.check(rc, "mm25_sum_parts")
    return da, db, dscale


def hopfield_split_eligible(M: int, N: int, d: int) -> bool:
    return bool(_lib.lib().mm25_hopfield_split_eligible(M, N, d))


@torch.library.custom_op("mm25::hopfield_bwd_split", mutates_args=(), device_types="cuda")
def hopfield_bwd_split(q4: Tensor, k4: Tensor, q_hi: Tensor, q_lo: Tensor, k_hi: Tensor,
                       k_lo: Tensor, d_o: Tensor, beta: Tensor, lse: Tensor, delta: Tensor,
                       need_dq: bool, need_dk: bool) -> Tuple[Tensor, Tensor]:
    """fp32-grade Hopfield backward on the split operands (include/mm25.h
    mm25_hopfield_bwd_split); q4 [M, W d], k4 [N, W d] bf16, d_o [M, d] fp32."""
    M, d4 = q4.shape
    N = k4.shape[0]
    d = d4 // int(_lib.lib().mm25_split_width())
    dev = q4.device
    L = _lib.lib()
    d_o = d_o.detach().float().contiguous()
    with torch.cuda.device(dev):
        dq = torch.empty((M, d) if need_dq else (0,), dtype=torch.float32, device=dev)
        dk = torch.empty((N, d) if need_dk else (0,), dtype=torch.float32, device=dev)
        ws = torch.empty((max(int(L.mm25_hopfield_bwd_split_workspace_bytes(M, N, d)), 16),),
                         dtype=torch.uint8, device=dev)
        rc = L.mm25_hopfield_bwd_split(_p(q4), _p(k4), _p(q_hi), _p(q_lo), _p(k_hi), _p(k_lo),
                                       _p(d_o), M, N, d, _p(beta), _p(lse.contiguous()),
                                       _p(delta.contiguous()), _p(dq) if need_dq else None,
                                       _p(dk) if need_dk else None, _p(ws), ws.numel(),
                                       _stream(q4))
    _lib.check(rc, "mm25_hopfield_bwd_split")
    return dq, dk


@torch.library.custom_op("mm25::col_merge", mutates_args=(), device_types="cuda")
def col_merge(col_max: Tensor, col_sum: Tensor) -> Tensor:
    """Merge [P, N] partial column (max, sum) in fixed order into col_lse [N]."""
    col_max = col_max.contiguous()
    col_sum = col_sum.contiguous()
    P, N = col_max.shape
    with torch.cuda.device(col_max.device):
        out = torch.empty((N,), dtype=torch.float32, device=col_max.device)
        rc = _lib.lib().mm25_col_merge(_p(col_max), _p(col_sum), P, N, _p(out),
                                       _stream(col_max))
    _lib.check(rc, "mm25_col_merge")
    return out


@torch.library.custom_op("mm25::clip_loss_finalize", mutates_args=(), device_types="cuda")
def clip_loss_finalize(row_lse: Tensor, diag: Tensor, col_lse: Optional[Tensor],
                       sx: Optional[Tensor], sy: Optional[Tensor], n_norm: int, scale: Tensor,
                       scale_kind: int, clamp_max: float, eps_ls: float) -> Tensor:
    """Closed-form CLIP loss per batch entry from the statistics."""
    row_lse = row_lse.contiguous()
    B, M = row_lse.shape
    N = M if col_lse is None else col_lse.shape[-1]
    d = 0 if sx is None else sx.shape[-1]
    with torch.cuda.device(row_lse.device):
        loss = torch.empty((B,), dtype=torch.float32, device=row_lse.device)
        rc = _lib.lib().mm25_clip_loss_finalize(_p(row_lse), _p(diag), _p(col_lse), _p(sx),
                                                _p(sy), B, M, N, d, n_norm, _p(scale),
                                                scale_kind, clamp_max, eps_ls, _p(loss),
                                                _stream(row_lse))
    _lib.check(rc, "mm25_clip_loss_finalize")
    return loss


@torch.library.custom_op("mm25::rowdot", mutates_args=(), device_types="cuda")
def rowdot(x: Tensor, y: Tensor) -> Tensor:
    """out[i] = x_i . y_i for fp32 [rows, d] matrices."""
    x = x.contiguous()
    y = y.contiguous()
    rows, d = x.shape
    with torch.cuda.device(x.device):
        out = torch.empty((rows,), dtype=torch.float32, device=x.device)
        rc = _lib.lib().mm25_rowdot(_p(x), _p(y), _p(out), rows, d, _stream(x))
    _lib.check(rc, "mm25_rowdot")
    return out


@torch.library.custom_op("mm25::hopfield_bwd_pass", mutates_args=(), device_types="cuda")
def hopfield_bwd_pass(a: Tensor, b: Tensor, a2: Tensor, b2: Tensor, bm1: Tensor,
                      bm2: Optional[Tensor], beta: float, lse: Tensor, D: Tensor,
                      stat_by_col: bool) -> Tensor:
    """One row-stationary pass of the Hopfield-retrieval backward (fp32, see include/mm25.h)."""
    a, b, a2, b2, bm1 = (t.contiguous() for t in (a, b, a2, b2, bm1))
    M, d = a.shape
    N = b.shape[0]
    with torch.cuda.device(a.device):
        out = torch.empty((M, d), dtype=torch.float32, device=a.device)
        rc = _lib.lib().mm25_hopfield_bwd_pass(_p(a), _p(b), _p(a2), _p(b2), _p(bm1),
                                               _p(bm2.contiguous()) if bm2 is not None else None,
                                               M, N, d, beta, _p(lse), _p(D), int(stat_by_col),
                                               _p(out), _stream(a))
    _lib.check(rc, "mm25_hopfield_bwd_pass")
    return out


def hopfield_tc_eligible(q: Tensor, k: Tensor) -> bool:
    """bf16 retrieval shapes the tensor-core Hopfield kernels take (include/mm25.h)."""
    if q.dtype != torch.bfloat16 or k.dtype != torch.bfloat16 or not q.is_cuda:
        return False
    M, d = q.shape[-2], q.shape[-1]
    N = k.shape[-2]
    return int(_lib.lib().mm25_hopfield_bwd_tc_workspace_bytes(M, N, d)) > 0


@torch.library.custom_op("mm25::hopfield_bwd_tc", mutates_args=(), device_types="cuda")
def hopfield_bwd_tc(q: Tensor, k: Tensor, d_o: Tensor, beta: Tensor, lse: Tensor, delta: Tensor,
                    need_dq: bool, need_dk: bool) -> Tuple[Tensor, Tensor]:
    """Hopfield-retrieval backward on the tensor cores: (dq [M,d], dk [N,d]) fp32 from bf16 q, k,
    fp32 dL/dO, the forward's row LSE and delta_i = dO_i . O_i (include/mm25.h)."""
    q, k, d_o = q.contiguous(), k.contiguous(), d_o.contiguous()
    M, d = q.shape
    N = k.shape[0]
    dev = q.device
    L = _lib.lib()
    with torch.cuda.device(dev):
        dq = torch.empty((M, d) if need_dq else (0,), dtype=torch.float32, device=dev)
        dk = torch.empty((N, d) if need_dk else (0,), dtype=torch.float32, device=dev)
        wsb = int(L.mm25_hopfield_bwd_tc_workspace_bytes(M, N, d))
        if wsb == 0:
            raise _lib.MM25Error("mm25_hopfield_bwd_tc: shape not eligible (M %% 128, N %% 256, "
                                 "d %% 256, d <= 1024): M=%d N=%d d=%d" % (M, N, d))
        ws = torch.empty((wsb,), dtype=torch.uint8, device=dev)
        rc = L.mm25_hopfield_bwd_tc(_p(q), _p(k), _p(d_o), M, N, d, _p(beta), _p(lse), _p(delta),
                                    _p(dq) if need_dq else None, _p(dk) if need_dk else None,
                                    _p(ws), ws.numel(), _stream(q))
    _lib.check(rc, "mm25_hopfield_bwd_tc")
    return dq, dk


def scale_param_grad(dscale: Tensor, scale: Tensor, scale_kind: int, clamp_max: float) -> Tensor:
    """Chain dL/ds to the raw scale parameter (clamp passes gradient at equality); one kernel."""
    s32 = scale.detach().reshape(1).to(torch.float32).contiguous()
    with torch.cuda.device(dscale.device):
        out = torch.empty((1,), dtype=torch.float32, device=dscale.device)
        rc = _lib.lib().mm25_scale_param_grad(_p(dscale), _p(s32), scale_kind, clamp_max, _p(out),
                                              _stream(dscale))
    _lib.check(rc, "mm25_scale_param_grad")
    return out.reshape(scale.shape).to(scale.dtype)


def peer_bcast(src: Tensor, peer_ptrs: List[int], dst_offset_bytes: int) -> None:
    """One-shot all-gather piece: store the contiguous `src` at `dst_offset_bytes` of every
    peer-mapped buffer in `peer_ptrs` (include/mm25.h mm25_peer_bcast).  Barrier afterwards."""
    import ctypes
    src = src.contiguous()
    P = len(peer_ptrs)
    table = (ctypes.c_void_p * P)(*peer_ptrs)
    with torch.cuda.device(src.device):
        rc = _lib.lib().mm25_peer_bcast(_p(src), src.numel() * src.element_size(), table, P,
                                        dst_offset_bytes, _stream(src))
    _lib.check(rc, "mm25_peer_bcast")


def sharded_loss(payload: Tensor, orow: int, col_lse: Tensor, ox: int, oy: int, d: int,
                 scale: Tensor, scale_kind: int, clamp_max: float, coef: float) -> Tensor:
    """Scalar tail of the row-sharded forward in one launch (include/mm25.h mm25_sharded_loss)."""
    P, ld = payload.shape
    col_lse = col_lse.contiguous()
    with torch.cuda.device(payload.device):
        loss = torch.empty((1,), dtype=torch.float32, device=payload.device)
        rc = _lib.lib().mm25_sharded_loss(_p(payload), P, ld, orow, _p(col_lse), col_lse.numel(),
                                          ox, oy, d, _p(scale), scale_kind, clamp_max, coef,
                                          _p(loss), _stream(payload))
    _lib.check(rc, "mm25_sharded_loss")
    return loss


def peer_bcast_mc(src: Tensor, mc_ptr: int, dst_offset_bytes: int) -> None:
    """peer_bcast through the NVSwitch multicast mapping of the symmetric buffer: one
    `multimem.st` reaches every rank's copy (include/mm25.h mm25_peer_bcast_mc)."""
    import ctypes
    src = src.contiguous()
    with torch.cuda.device(src.device):
        rc = _lib.lib().mm25_peer_bcast_mc(_p(src), src.numel() * src.element_size(),
                                           ctypes.c_void_p(int(mc_ptr)), dst_offset_bytes,
                                           _stream(src))
    _lib.check(rc, "mm25_peer_bcast_mc")


def col_merge_ld(payload: Tensor, N: int) -> Tensor:
    """col_merge on a gathered [P, ld] payload whose rows start with (col_max[N], col_sum[N])."""
    P, ld = payload.shape
    with torch.cuda.device(payload.device):
        out = torch.empty((N,), dtype=torch.float32, device=payload.device)
        rc = _lib.lib().mm25_col_merge_ld(_p(payload), payload.data_ptr() + 4 * N, P, N, ld,
                                          _p(out), _stream(payload))
    _lib.check(rc, "mm25_col_merge_ld")
    return out


def smooth_term_(loss: Tensor, payload: Tensor, ox: int, oy: int, d: int, scale: Tensor,
                 scale_kind: int, clamp_max: float, coef: float) -> Tensor:
    """loss[0] -= coef * s * (sum_p payload[p, ox:ox+d]) . (sum_p payload[p, oy:oy+d]) in place
    (label-smoothing term of the sharded loss, include/mm25.h mm25_smooth_term)."""
    P, ld = payload.shape
    with torch.cuda.device(loss.device):
        rc = _lib.lib().mm25_smooth_term(_p(payload), P, ld, ox, oy, d, _p(scale), scale_kind,
                                         clamp_max, coef, _p(loss), _stream(loss))
    _lib.check(rc, "mm25_smooth_term")
    return loss


# ------------------------------------------------------------------------------------------------
# vCLIP sampling path (include/mm25.h "vCLIP sampling path").  All fp32, sample-major [S, B, d].
# ------------------------------------------------------------------------------------------------
VMF_CFG = (1e-10, 0.0, 1e-10, 0.0, -1.0)      # VonMisesFisher.py:181-183,224-236
POWER_SPHERICAL_CFG = (0.0, 1e-7, 0.0, 1e-7, 1e-7)  # power_spherical _EPS (unverified)


def _f32dev(t: Tensor, name: str) -> Tensor:
    if not t.is_cuda:
        raise _lib.MM25Error("%s must be a CUDA tensor: this path has no CPU fallback" % name)
    return t.detach().to(torch.float32).contiguous()


@torch.library.custom_op("mm25::projnormal_fwd", mutates_args=(), device_types="cuda")
def projnormal_fwd(mu: Tensor, sigma: Tensor, eps_in: Optional[Tensor], seed: int,
                   S: int) -> Tuple[Tensor, Tensor]:
    """ProjectedNormal.rsample((S,)) (ProjectedNormal.py:41-49) -> (samples [S,B,d], 1/norm [S,B])."""
    mu, sigma = _f32dev(mu, "mu"), _f32dev(sigma, "sigma")
    B, d = mu.shape
    eps = None if eps_in is None else _f32dev(eps_in, "eps_in")
    with torch.cuda.device(mu.device):
        out = torch.empty((S, B, d), dtype=torch.float32, device=mu.device)
        inv = torch.empty((S, B), dtype=torch.float32, device=mu.device)
        rc = _lib.lib().mm25_sample_projnormal_fwd(_p(mu), _p(sigma), _p(eps), seed, S, B, d,
                                                   _p(out), _p(inv), _stream(mu))
    _lib.check(rc, "mm25_sample_projnormal_fwd")
    return out, inv


@torch.library.custom_op("mm25::projnormal_bwd", mutates_args=(), device_types="cuda")
def projnormal_bwd(g: Tensor, y: Tensor, inv_norm: Tensor, eps_in: Optional[Tensor],
                   seed: int) -> Tuple[Tensor, Tensor]:
    g, y = _f32dev(g, "g"), _f32dev(y, "y")
    S, B, d = y.shape
    eps = None if eps_in is None else _f32dev(eps_in, "eps_in")
    with torch.cuda.device(y.device):
        dmu = torch.empty((B, d), dtype=torch.float32, device=y.device)
        dsg = torch.empty((B, d), dtype=torch.float32, device=y.device)
        rc = _lib.lib().mm25_sample_projnormal_bwd(_p(g), _p(y), _p(inv_norm), _p(eps), seed, S,
                                                   B, d, _p(dmu), _p(dsg), _stream(y))
    _lib.check(rc, "mm25_sample_projnormal_bwd")
    return dmu, dsg


@torch.library.custom_op("mm25::sphere_fwd", mutates_args=(), device_types="cuda")
def sphere_fwd(t: Tensor, loc: Tensor, v_in: Optional[Tensor], seed: int, v_clamp: float,
               v_add: float, h_clamp: float, h_add: float, omt2_min: float) -> Tensor:
    """[t, sqrt(1-t^2) v/|v|] reflected so that e1 -> loc; t [S,B], loc [B,d] -> [S,B,d]."""
    t, loc = _f32dev(t, "t"), _f32dev(loc, "loc")
    S, B = t.shape
    d = loc.shape[-1]
    v = None if v_in is None else _f32dev(v_in, "v_in")
    with torch.cuda.device(loc.device):
        out = torch.empty((S, B, d), dtype=torch.float32, device=loc.device)
        rc = _lib.lib().mm25_sphere_sample_fwd(_p(t), _p(loc), _p(v), seed, S, B, d, v_clamp,
                                               v_add, h_clamp, h_add, omt2_min, _p(out),
                                               _stream(loc))
    _lib.check(rc, "mm25_sphere_sample_fwd")
    return out


@torch.library.custom_op("mm25::sphere_bwd", mutates_args=(), device_types="cuda")
def sphere_bwd(g: Tensor, y: Tensor, t: Tensor, loc: Tensor, v_clamp: float, v_add: float,
               h_clamp: float, h_add: float, omt2_min: float) -> Tuple[Tensor, Tensor]:
    g, y, t, loc = _f32dev(g, "g"), _f32dev(y, "y"), _f32dev(t, "t"), _f32dev(loc, "loc")
    S, B, d = y.shape
    with torch.cuda.device(y.device):
        dloc = torch.empty((B, d), dtype=torch.float32, device=y.device)
        dt = torch.empty((S, B), dtype=torch.float32, device=y.device)
        rc = _lib.lib().mm25_sphere_sample_bwd(_p(g), _p(y), _p(t), _p(loc), S, B, d, v_clamp,
                                               v_add, h_clamp, h_add, omt2_min, _p(dloc), _p(dt),
                                               _stream(y))
    _lib.check(rc, "mm25_sphere_sample_bwd")
    return dloc, dt


@torch.library.custom_op("mm25::vmf_sample_w", mutates_args=(), device_types="cuda")
def vmf_sample_w(kappa: Tensor, d: int, S: int, seed: int, prop_eps: Optional[Tensor],
                 prop_u: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    """Wood rejection sampler (VonMisesFisher.py:184-222) -> (w [S,B], accepted count [B])."""
    kappa = _f32dev(kappa, "kappa").reshape(-1)
    B = kappa.numel()
    pe = None if prop_eps is None else _f32dev(prop_eps, "prop_eps")
    pu = None if prop_u is None else _f32dev(prop_u, "prop_u")
    P = 0 if pe is None else pe.shape[-1]
    with torch.cuda.device(kappa.device):
        w = torch.empty((S, B), dtype=torch.float32, device=kappa.device)
        cnt = torch.full((B,), S, dtype=torch.int32, device=kappa.device)
        rc = _lib.lib().mm25_vmf_sample_w(_p(kappa), d, S, B, seed, _p(pe), _p(pu), P, _p(w),
                                          _p(cnt), _stream(kappa))
    _lib.check(rc, "mm25_vmf_sample_w")
    return w, cnt


@torch.library.custom_op("mm25::beta_sample", mutates_args=(), device_types="cuda")
def beta_sample(alpha: Tensor, beta: Tensor, S: int, seed: int) -> Tensor:
    alpha, beta = _f32dev(alpha, "alpha").reshape(-1), _f32dev(beta, "beta").reshape(-1)
    B = alpha.numel()
    with torch.cuda.device(alpha.device):
        z = torch.empty((S, B), dtype=torch.float32, device=alpha.device)
        rc = _lib.lib().mm25_beta_sample(_p(alpha), _p(beta), S, B, seed, _p(z), _stream(alpha))
    _lib.check(rc, "mm25_beta_sample")
    return z


@torch.library.custom_op("mm25::beta_grad", mutates_args=(), device_types="cuda")
def beta_grad(z: Tensor, alpha: Tensor, beta: Tensor) -> Tensor:
    """dz/dalpha of z ~ Beta(alpha, beta) (implicit reparameterisation, as torch's Beta.rsample)."""
    z = _f32dev(z, "z")
    alpha, beta = _f32dev(alpha, "alpha").reshape(-1), _f32dev(beta, "beta").reshape(-1)
    S, B = z.shape
    with torch.cuda.device(z.device):
        out = torch.empty((S, B), dtype=torch.float32, device=z.device)
        rc = _lib.lib().mm25_beta_grad(_p(z), _p(alpha), _p(beta), S, B, _p(out), _stream(z))
    _lib.check(rc, "mm25_beta_grad")
    return out


@torch.library.custom_op("mm25::log_bessel", mutates_args=(), device_types="cuda")
def log_bessel(kappa: Tensor, nu: float) -> Tuple[Tensor, Tensor]:
    """(log I_nu(kappa), log I_{nu+1}(kappa)) in fp64 (replaces the mpmath round trip)."""
    if not kappa.is_cuda:
        raise _lib.MM25Error("kappa must be a CUDA tensor: this path has no CPU fallback")
    k = kappa.detach().to(torch.float64).contiguous()
    with torch.cuda.device(k.device):
        o0 = torch.empty_like(k)
        o1 = torch.empty_like(k)
        rc = _lib.lib().mm25_log_bessel(_p(k), k.numel(), nu, _p(o0), _p(o1), _stream(k))
    _lib.check(rc, "mm25_log_bessel")
    return o0, o1


@torch.library.custom_op("mm25::topk_rows", mutates_args=(), device_types="cuda")
def topk_rows(scores: Tensor, k: int) -> Tuple[Tensor, Tensor]:
    """(indices int64 [Q,k], values [Q,k]) of the k largest entries per row of fp32 scores,
    sorted descending (metrics/topk.py:43-48); exact, ties lowest index first."""
    if not scores.is_cuda:
        raise _lib.MM25Error("scores must be a CUDA tensor: this path has no CPU fallback")
    s = scores.detach().to(torch.float32)
    if s.dim() == 1:
        s = s.unsqueeze(0)
    s = s.contiguous()
    Q, N = s.shape
    with torch.cuda.device(s.device):
        idx = torch.empty((Q, k), dtype=torch.int64, device=s.device)
        val = torch.empty((Q, k), dtype=torch.float32, device=s.device)
        L = _lib.lib()
        if os.environ.get("MM25_TOPK_PLAIN"):   # A/B switch: the one-CTA-per-row kernel
            rc = L.mm25_topk_rows(_p(s), Q, N, N, k, _p(idx), _p(val), _stream(s))
        else:
            wsb = int(L.mm25_topk_workspace_bytes(Q, N, k))
            ws = torch.empty((max(wsb, 16),), dtype=torch.uint8, device=s.device)
            rc = L.mm25_topk_rows_ws(_p(s), Q, N, N, k, _p(idx), _p(val), _p(ws), ws.numel(),
                                     _stream(s))
    _lib.check(rc, "mm25_topk_rows")
    return idx, val


@torch.library.custom_op("mm25::rank_both", mutates_args=("row_rank", "col_rank"),
                         device_types="cuda")
def rank_both(scores: Tensor, row0: int, row_tval: Optional[Tensor], row_tcol: Optional[Tensor],
              col_tval: Optional[Tensor], col_trow: Optional[Tensor],
              row_rank: Optional[Tensor], col_rank: Optional[Tensor]) -> None:
    """One pass over the fp32 slab `scores` [Q, N] (global rows row0..) that ADDS into
    row_rank[q] / col_rank[j] (int32) how many entries sort ahead of the row / column target in
    the stable descending order (include/mm25.h mm25_rank_both; train.py:338-356)."""
    if not scores.is_cuda:
        raise _lib.MM25Error("scores must be a CUDA tensor: this path has no CPU fallback")
    if scores.dtype != torch.float32 or scores.dim() != 2 or scores.stride(1) != 1:
        raise ValueError("scores must be a row-major fp32 [Q, N] slab")
    Q, N = scores.shape
    for name, t, n, dt in (("row_tval", row_tval, Q, torch.float32),
                           ("row_tcol", row_tcol, Q, torch.int64),
                           ("row_rank", row_rank, Q, torch.int32),
                           ("col_tval", col_tval, N, torch.float32),
                           ("col_trow", col_trow, N, torch.int64),
                           ("col_rank", col_rank, N, torch.int32)):
        if t is not None and (t.dtype != dt or t.numel() != n or not t.is_contiguous()
                              or t.device != scores.device):
            raise ValueError("%s must be a contiguous %s tensor of %d elements on %s"
                             % (name, dt, n, scores.device))
    with torch.cuda.device(scores.device):
        rc = _lib.lib().mm25_rank_both(_p(scores), Q, N, scores.stride(0), row0, _p(row_tval),
                                       _p(row_tcol), _p(col_tval), _p(col_trow), _p(row_rank),
                                       _p(col_rank), _stream(scores))
    _lib.check(rc, "mm25_rank_both")


@torch.library.custom_op("mm25::sim_rank", mutates_args=("rank",), device_types="cuda")
def sim_rank(a: Tensor, b: Tensor, tval: Tensor, tcol: Tensor, rank: Tensor) -> None:
    """Ranks fused into the similarity GEMM (include/mm25.h mm25_sim_rank): ADDS into rank[q]
    (int32) the number of gallery rows j != tcol[q] whose score a_q . b_j sorts ahead of the
    target (tval[q], tcol[q]) in the stable descending order; the [Q, N] scores are never
    written.  a [Q, d], b [N, d] bf16."""
    if not (a.is_cuda and b.is_cuda):
        raise _lib.MM25Error("sim_rank needs CUDA tensors: this path has no CPU fallback")
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16 or a.dim() != 2 or b.dim() != 2:
        raise TypeError("a, b must be 2-D bf16")
    a, b = a.contiguous(), b.contiguous()
    Q, d = a.shape
    N = b.shape[0]
    for name, t, dt in (("tval", tval, torch.float32), ("tcol", tcol, torch.int64),
                        ("rank", rank, torch.int32)):
        if t.dtype != dt or t.numel() != Q or not t.is_contiguous() or t.device != a.device:
            raise ValueError("%s must be a contiguous %s tensor of %d elements on %s"
                             % (name, dt, Q, a.device))
    with torch.cuda.device(a.device):
        rc = _lib.lib().mm25_sim_rank(_p(a), _p(b), BF16, Q, N, d, _p(tval), _p(tcol), _p(rank),
                                      _stream(a))
    _lib.check(rc, "mm25_sim_rank")


def sim_rank_eligible(a: Tensor, b: Tensor) -> bool:
    d = a.shape[-1]
    return (a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and
            d % 64 == 0 and 64 <= d <= 1024)


@torch.library.custom_op("mm25::sim_bwd_rs", mutates_args=(), device_types="cuda")
def sim_bwd_rs(a: Tensor, b: Tensor, scale: Tensor, scale_kind: int, clamp_max: float,
               diag_offset: int, row_lse: Tensor, col_lse: Tensor, w_row: float, w_col: float,
               w_diag: float, w_all: float, gout: Optional[Tensor], peer_ptrs: List[int],
               dot_bound: float, e_keep: Optional[Tensor] = None,
               e_flag: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """Fused backward + reduce-scatter (include/mm25.h mm25_sim_bwd_rs): returns (da, dscale);
    the partial db of this rank is added into the peers' [N/P, d] fp32 buffers `peer_ptrs`."""
    a = _cuda3(a, "a")
    b = _cuda3(b, "b")
    _, M, d = a.shape
    N = b.shape[1]
    dev = a.device
    L = _lib.lib()
    P = len(peer_ptrs)
    import ctypes
    table = (ctypes.c_void_p * P)(*peer_ptrs)
    with torch.cuda.device(dev):
        da = torch.empty((1, M, d), dtype=torch.float32, device=dev)
        nparts = int(L.mm25_sim_bwd_nparts(1, M, N, d, _dt(a), IMPL_TC))
        parts = torch.zeros((max(nparts, 1),), dtype=torch.float32, device=dev)
        dscale = torch.empty((1,), dtype=torch.float32, device=dev)
        with_e = e_keep is not None and e_keep.numel() > 0 and e_flag is not None
        wsb = L.mm25_sim_bwd_workspace_bytes(1, M, N, d, _dt(a), IMPL_TC, int(with_e))
        ws = torch.empty((max(int(wsb), 16),), dtype=torch.uint8, device=dev)
        rc = L.mm25_sim_bwd_rs(_p(a), _p(b), _dt(a), M, N, d, _p(scale), scale_kind, clamp_max,
                               diag_offset, _p(row_lse), _p(col_lse), w_row, w_col, w_diag, w_all,
                               _p(gout), _p(da), table, P, _p(parts), _p(ws), ws.numel(),
                               dot_bound, _p(e_keep) if with_e else None,
                               _p(e_flag) if with_e else None, _stream(a))
        _lib.check(rc, "mm25_sim_bwd_rs")
        rc = L.mm25_sum_parts(_p(parts), parts.numel(), 1.0, _p(dscale), _stream(a))
    _lib.check(rc, "mm25_sum_parts")
    return da, dscale


def tc_eligible_bwd(a: Tensor, b: Tensor) -> bool:
    """True when the tcgen05 backward (and so the fused reduce-scatter) takes these operands
    (sim_tc_bwd.cu eligible_bwd: bf16, d % 256 == 0, d <= 1024, M % 128 == 0, N % 256 == 0)."""
    M, d = a.shape[-2], a.shape[-1]
    N = b.shape[-2]
    return (a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and d % 256 == 0 and
            d <= 1024 and M % 128 == 0 and N % 256 == 0)

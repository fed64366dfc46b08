"""Retrieval metrics of the AlignCLIP trainer without the N x N sort.

Reference: `AlignCLIP/main/train.py:338-356 get_clip_metrics` (argsort of both logit matrices on
the CPU, then the position of the ground truth) and `AlignCLIP/main/retrieval.py:55-102
recall_at_k` (the same argsort, then `topk == map` comparisons per k).  The position of a target
in a descending sort is the number of entries that sort ahead of it, so one HBM-bound pass of
`mm25_rank_both` over the scores gives every rank of BOTH directions (rows and columns) — no
sort, no transpose, no device->host copy of the matrix.  Ranks are integers: identical to the
reference's whenever the scores are tie-free (torch.argsort leaves tie order unspecified; here
ties break lowest index first, the stable order).

The scores themselves are computed exactly as the reference writes them
(`logit_scale * image_features @ text_features.t()`, `text_encodings @ image_encodings.T`), a
plain library GEMM in the features' dtype.  When the matrix does not fit `max_score_bytes` it is
produced in row slabs, twice (first pass: the diagonal = column targets; second pass: ranks).
"""
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch import Tensor

from .. import ops as _ops  # tests swap this for the dense CPU stand-in (tests/cpu_backend.py)

# N x N fp32 at N = 65536 is 16 GiB of the 180 GB: materialise up to this, slab above
MAX_SCORE_BYTES = 32 << 30
_MAX_ROWS_PER_CALL = 65535 * 32  # grid.y limit of mm25_rank_both


def _zeros_i32(n: int, like: Tensor) -> Tensor:
    return torch.zeros((n,), dtype=torch.int32, device=like.device)


def _rank_call(scores: Tensor, row0: int, row_t, col_t, row_rank, col_rank) -> None:
    """mm25_rank_both over `scores`, split so that no call exceeds the grid limit."""
    s = scores if scores.dtype == torch.float32 else scores.float()
    if s.stride(-1) != 1:
        s = s.contiguous()
    Q = s.shape[0]
    for a in range(0, Q, _MAX_ROWS_PER_CALL):
        b = min(Q, a + _MAX_ROWS_PER_CALL)
        rt = None if row_t is None else (row_t[0][a:b], row_t[1][a:b])
        _ops.rank_both(s[a:b], row0 + a,
                       None if rt is None else rt[0], None if rt is None else rt[1],
                       None if col_t is None else col_t[0], None if col_t is None else col_t[1],
                       None if row_rank is None else row_rank[a:b], col_rank)


def retrieval_ranks(scores: Tensor, row_targets: Optional[Tensor] = None,
                    col_targets: Optional[Tensor] = None) -> Tuple[Optional[Tensor],
                                                                   Optional[Tensor]]:
    """0-based positions of `scores[q, row_targets[q]]` in the descending order of row q and of
    `scores[col_targets[j], j]` in the descending order of column j (int32 tensors; None for a
    direction whose targets are None).  Equals `where(argsort(scores, descending=True) ==
    target)[1]` of train.py:349-350 for rows, and the same on `scores.t()` for columns."""
    if scores.dim() != 2:
        raise ValueError("scores must be [Q, N]")
    s = scores.detach()
    s = s if s.dtype == torch.float32 else s.float()
    Q, N = s.shape
    row_t = col_t = row_rank = col_rank = None
    if row_targets is not None:
        rt = row_targets.to(device=s.device, dtype=torch.int64).contiguous()
        row_t = (s.gather(1, rt.view(-1, 1)).view(-1).contiguous(), rt)
        row_rank = _zeros_i32(Q, s)
    if col_targets is not None:
        ct = col_targets.to(device=s.device, dtype=torch.int64).contiguous()
        col_t = (s.gather(0, ct.view(1, -1)).view(-1).contiguous(), ct)
        col_rank = _zeros_i32(N, s)
    _rank_call(s, 0, row_t, col_t, row_rank, col_rank)
    return row_rank, col_rank


def _scores(logit_scale, a: Tensor, b: Tensor) -> Tensor:
    # the reference's expression, evaluated left to right: (logit_scale * a) @ b.t()
    return (logit_scale * a @ b.t()) if logit_scale is not None else a @ b.t()


def fused_ranks(queries: Tensor, gallery: Tensor, targets: Tensor) -> Tensor:
    """0-based rank of gallery row `targets[q]` among the scores `queries[q] . gallery[j]`, with
    the similarity GEMM and the count fused in one tcgen05 kernel (`mm25_sim_rank`): the
    [Q, N] score matrix is never formed.  bf16 features (fp32 accumulation); the target's score
    is an fp32 dot product of the same bf16 rows."""
    q = queries.detach().to(torch.bfloat16).contiguous()
    g = gallery.detach().to(torch.bfloat16).contiguous()
    t = targets.to(device=q.device, dtype=torch.int64).contiguous()
    tval = (q.float() * g[t].float()).sum(dim=-1).contiguous()
    rank = _zeros_i32(q.shape[0], q)
    _ops.sim_rank(q, g, tval, t, rank)
    return rank


def _use_fused(fused, x: Tensor, y: Tensor, logit_scale=None) -> bool:
    """fused=None: the exact route (library GEMM in the reference's expression +
    `mm25_rank_both` on identical scores) unless the environment sets MM25_FUSED_RANKS=1; the
    fused kernel ranks fp32-accumulated dot products of the bf16 rows, which is not the order of
    the reference's own bf16-ROUNDED scores wherever the rounding creates ties, so it is an
    explicit choice: `fused=True` (any CUDA features of an eligible width; fp32 features are
    rounded to bf16 first) or the environment switch."""
    ok = (hasattr(_ops, "sim_rank") and x.is_cuda and y.is_cuda and x.shape[-1] % 64 == 0
          and 64 <= x.shape[-1] <= 1024 and x.shape[-1] == y.shape[-1])
    if ok and logit_scale is not None and float(logit_scale) <= 0:
        ok = False  # a non-positive scale changes the order: rank the scaled scores instead
    if fused is None:
        import os
        return bool(ok and os.environ.get("MM25_FUSED_RANKS") == "1"
                    and x.dtype == torch.bfloat16 and y.dtype == torch.bfloat16)
    if fused and not ok:
        raise ValueError("fused ranks need CUDA features with d % 64 == 0, 64 <= d <= 1024 and "
                         "a positive logit scale")
    return bool(fused)


def paired_ranks(image_features: Tensor, text_features: Tensor, logit_scale=None,
                 max_score_bytes: Optional[int] = None, fused=None) -> Tuple[Tensor, Tensor]:
    """(image_to_text, text_to_image) 0-based ranks of the paired item i <-> i (int32 [N])."""
    x, y = image_features.detach(), text_features.detach()
    if x.shape[0] != y.shape[0]:
        raise ValueError("paired ranks need as many images as texts")
    N = x.shape[0]
    if _use_fused(fused, x, y, logit_scale):
        ar = torch.arange(N, device=x.device, dtype=torch.int64)
        return fused_ranks(x, y, ar), fused_ranks(y, x, ar)
    budget = MAX_SCORE_BYTES if max_score_bytes is None else int(max_score_bytes)
    ar = torch.arange(N, device=x.device, dtype=torch.int64)
    if N * N * 4 <= budget:
        return retrieval_ranks(_scores(logit_scale, x, y), ar, ar)
    rows = max(32, min(N, (budget // (4 * N)) // 32 * 32))
    diag = torch.empty((N,), dtype=torch.float32, device=x.device)
    for a in range(0, N, rows):  # pass 1: the targets (bit-identical to what pass 2 recomputes)
        b = min(N, a + rows)
        diag[a:b] = _scores(logit_scale, x[a:b], y).float().diagonal(offset=a)
    i2t, t2i = _zeros_i32(N, x), _zeros_i32(N, x)
    for a in range(0, N, rows):  # pass 2: every score feeds its row count and its column count
        b = min(N, a + rows)
        _rank_call(_scores(logit_scale, x[a:b], y), a, (diag[a:b].contiguous(), ar[a:b]),
                   (diag, ar), i2t[a:b], t2i)
    return i2t, t2i


def get_clip_metrics(image_features: Tensor, text_features: Tensor, logit_scale,
                     max_score_bytes: Optional[int] = None, fused=None) -> dict:
    """Drop-in for train.py:338-356: mean / median rank and R@{1,5,10} of both directions, same
    keys, same numpy arithmetic on the ranks; only 2 N int32 ranks leave the device."""
    i2t, t2i = paired_ranks(image_features, text_features, logit_scale, max_score_bytes, fused)
    metrics = {}
    for name, rank in (("image_to_text", i2t), ("text_to_image", t2i)):
        preds = rank.cpu().numpy().astype(np.int64)
        metrics[f"{name}_mean_rank"] = preds.mean() + 1
        metrics[f"{name}_median_rank"] = np.floor(np.median(preds)) + 1
        for k in [1, 5, 10]:
            metrics[f"{name}_R@{k}"] = np.mean(preds < k)
    return metrics


def recall_at_k_from_encodings(image_encodings: Tensor, text_encodings: Tensor,
                               text_to_image_map: Tensor, image_to_text_map: Tensor,
                               k_vals: Sequence[int], fused=None) -> Tuple[List[float],
                                                                           List[float]]:
    """retrieval.py:59-102 after `encode_dataset`: (text_to_image_recall, image_to_text_recall),
    one entry per k.  `topk[:, :k]` contains the target iff its rank is < k, so one rank per text
    (row direction) and one per caption (column direction, one pass per caption slot) replace
    the two argsorts."""
    num_text = text_encodings.shape[0]
    num_im = image_encodings.shape[0]
    captions_per_image = image_to_text_map.shape[1]
    if captions_per_image > 0 and _use_fused(fused, text_encodings, image_encodings):
        # fused GEMM + count: text -> image once; image -> text once per caption slot (the best
        # caption's rank decides `topk[:, :k] == caption` for any k)
        dev = text_encodings.device
        text_rank = fused_ranks(text_encodings, image_encodings, text_to_image_map)
        i2t_map = image_to_text_map.to(device=dev, dtype=torch.int64)
        best = None
        for c in range(captions_per_image):
            cap_rank = fused_ranks(image_encodings, text_encodings, i2t_map[:, c])
            best = cap_rank if best is None else torch.minimum(best, cap_rank)
        ks = torch.tensor(list(k_vals), dtype=torch.int32, device=dev).view(-1, 1)
        t_correct = (text_rank.view(1, -1) < ks).sum(dim=1).tolist()
        i_correct = (best.view(1, -1) < ks).sum(dim=1).tolist()
        return ([c / num_text for c in t_correct], [c / num_im for c in i_correct])
    dist_matrix = (text_encodings.detach() @ image_encodings.detach().T).float()
    dev = dist_matrix.device
    t2i_map = text_to_image_map.to(device=dev, dtype=torch.int64).contiguous()
    i2t_map = image_to_text_map.to(device=dev, dtype=torch.int64)

    text_rank = _zeros_i32(num_text, dist_matrix)
    row_t = (dist_matrix.gather(1, t2i_map.view(-1, 1)).view(-1).contiguous(), t2i_map)
    best = None
    for c in range(captions_per_image):
        cap = i2t_map[:, c].contiguous()
        col_t = (dist_matrix.gather(0, cap.view(1, -1)).view(-1).contiguous(), cap)
        cap_rank = _zeros_i32(num_im, dist_matrix)
        _rank_call(dist_matrix, 0, row_t if c == 0 else None, col_t,
                   text_rank if c == 0 else None, cap_rank)
        best = cap_rank if best is None else torch.minimum(best, cap_rank)
    if captions_per_image == 0:
        _rank_call(dist_matrix, 0, row_t, None, text_rank, None)
        best = torch.full((num_im,), num_text, dtype=torch.int32, device=dev)
    ks = torch.tensor(list(k_vals), dtype=torch.int32, device=dev).view(-1, 1)
    t_correct = (text_rank.view(1, -1) < ks).sum(dim=1).tolist()  # one D2H per direction
    i_correct = (best.view(1, -1) < ks).sum(dim=1).tolist()
    return ([c / num_text for c in t_correct], [c / num_im for c in i_correct])


def recall_at_k(clip, dataset, device, k_vals: List[int], batch_size: int, encode_dataset=None):
    """Signature of retrieval.py:55.  The encoder pass (`encode_dataset`, retrieval.py:7-52) is
    outside this path: pass the reference's function (or any callable with its signature and
    its 4 return values) as `encode_dataset`."""
    if encode_dataset is None:
        raise ValueError("recall_at_k needs the caller's encode_dataset(clip, dataset, device, "
                         "batch_size=...) (AlignCLIP/main/retrieval.py:7-52); the encoders are "
                         "not part of this library")
    image_encodings, text_encodings, text_to_image_map, image_to_text_map = encode_dataset(
        clip, dataset, device, batch_size=batch_size)
    return recall_at_k_from_encodings(image_encodings, text_encodings, text_to_image_map,
                                      image_to_text_map, k_vals)

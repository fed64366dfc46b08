"""GPU: retrieval ranks fused into the similarity GEMM (`mm25_sim_rank`, csrc/sim_tc_rank.cu;
SURVEY §8f rank 1 — AlignCLIP/main/train.py:338-356, main/retrieval.py:55-102,
metrics/topk.py:54-82 without the Q x N score matrix).

The kernel ranks fp32-accumulated dot products of bf16 rows.  A materialised fp32 matmul of the
same bf16 rows differs from the tensor-core accumulators only in summation order (<= ~1e-6 on
unit vectors), so the exact statement is a bracket: with delta = 5e-6,
    #{z_qj > t_q + delta}  <=  rank[q]  <=  #{z_qj >= t_q - delta}   (j != target),
and equality with the sort-based oracle wherever no score lies within delta of the target.
On exactly representable scores (small-integer features) the result is bit-exact, ties included.
"""
import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO

pytestmark = pytest.mark.gpu

DELTA = 5e-6


def _bracket_check(q, g, tgt, rank):
    z = q.float() @ g.float().T
    t = z.gather(1, tgt.view(-1, 1))
    col = torch.arange(g.shape[0], device=q.device).view(1, -1)
    other = col != tgt.view(-1, 1)
    lo = ((z > t + DELTA) & other).sum(dim=1)
    hi = ((z >= t - DELTA) & other).sum(dim=1)
    r = rank.long()
    assert bool(((r >= lo) & (r <= hi)).all()), (int((r < lo).sum()), int((r > hi).sum()))
    clear = lo == hi                      # no near-tie: must equal the stable-sort rank
    want = MO.ranks_by_sort(z.cpu(), tgt.cpu())
    assert torch.equal(r.cpu()[clear.cpu()], want[clear.cpu()])
    return float(clear.float().mean())


@pytest.mark.parametrize("Q,N,d", [(128, 256, 64), (1, 1, 64), (5, 7, 128), (130, 1000, 64),
                                   (300, 4099, 256), (2048, 2048, 512), (777, 33000, 1024)])
def test_sim_rank_bracket_vs_materialised_scores(lib_built, Q, N, d):
    from multimodal_2025_b200.metrics.retrieval import fused_ranks
    g = torch.Generator().manual_seed(Q * 7 + N)
    q = torch.nn.functional.normalize(torch.randn(Q, d, generator=g), dim=-1).bfloat16().cuda()
    gal = torch.nn.functional.normalize(torch.randn(N, d, generator=g), dim=-1).bfloat16().cuda()
    tgt = torch.randint(0, N, (Q,), generator=g).cuda()
    rank = fused_ranks(q, gal, tgt)
    assert rank.dtype == torch.int32 and rank.shape == (Q,)
    frac_clear = _bracket_check(q, gal, tgt, rank)
    assert frac_clear > (0.5 if N <= 4099 else 0.0)   # rows without a near-tie: checked exactly


@pytest.mark.parametrize("Q,N,d", [(100, 600, 64), (257, 2049, 128)])
def test_sim_rank_exact_on_integer_features_with_ties(lib_built, Q, N, d):
    """features in {-1, 0, 1}: every dot product is an exactly representable integer, so the
    accumulators equal the materialised scores bit for bit and most entries TIE with the target;
    ranks must equal the stable descending sort (lowest column first among equals)."""
    from multimodal_2025_b200.metrics.retrieval import fused_ranks
    g = torch.Generator().manual_seed(N)
    q = torch.randint(-1, 2, (Q, d), generator=g).float()
    gal = torch.randint(-1, 2, (N, d), generator=g).float()
    tgt = torch.randint(0, N, (Q,), generator=g)
    want = MO.ranks_by_sort(q @ gal.T, tgt)
    got = fused_ranks(q.bfloat16().cuda(), gal.bfloat16().cuda(), tgt.cuda())
    assert torch.equal(got.cpu().long(), want)


def test_sim_rank_accumulates_and_is_repeatable(lib_built):
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(3)
    q = torch.randn(256, 128, generator=g).bfloat16().cuda()
    gal = torch.randn(4096, 128, generator=g).bfloat16().cuda()
    tgt = torch.randint(0, 4096, (256,), generator=g).cuda()
    tval = (q.float() * gal[tgt].float()).sum(-1).contiguous()
    r1 = torch.zeros(256, dtype=torch.int32, device="cuda")
    ops.sim_rank(q, gal, tval, tgt, r1)
    r2 = r1.clone()
    ops.sim_rank(q, gal, tval, tgt, r2)          # ADDS: second call doubles
    assert torch.equal(r2, 2 * r1)


def test_get_clip_metrics_fused_vs_unfused_and_oracle(lib_built):
    """bf16 features, fused=True: the metrics agree with the unfused route
    (library GEMM + mm25_rank_both on fp32 features of the same bf16 values) and the oracle."""
    from multimodal_2025_b200.metrics import get_clip_metrics
    from multimodal_2025_b200.metrics.retrieval import paired_ranks
    N, d = 3000, 256
    g = torch.Generator().manual_seed(1)
    x = torch.nn.functional.normalize(torch.randn(N, d, generator=g), dim=-1)
    y = torch.nn.functional.normalize(x + 0.9 * torch.randn(N, d, generator=g), dim=-1)
    xb, yb = x.bfloat16().cuda(), y.bfloat16().cuda()
    fused = get_clip_metrics(xb, yb, torch.tensor(14.3, device="cuda"), fused=True)
    plain = get_clip_metrics(xb.float(), yb.float(), torch.tensor(14.3, device="cuda"),
                             fused=False)
    want = MO.get_clip_metrics(xb.float().cpu(), yb.float().cpu(), torch.tensor(14.3))
    for k in want:
        assert abs(float(fused[k]) - float(want[k])) <= 2e-3 * max(1.0, abs(float(want[k]))), k
        assert abs(float(plain[k]) - float(want[k])) <= 2e-3 * max(1.0, abs(float(want[k]))), k
    i2t, t2i = paired_ranks(xb, yb, fused=True)
    j2t, u2i = paired_ranks(xb.float(), yb.float(), fused=False)
    assert float((i2t != j2t).float().mean()) < 5e-3 and float((t2i != u2i).float().mean()) < 5e-3
    assert int((i2t - j2t).abs().max()) <= 2 and int((t2i - u2i).abs().max()) <= 2


def test_recall_at_k_and_top_k_score_fused(lib_built):
    from multimodal_2025_b200.metrics.retrieval import recall_at_k_from_encodings
    from multimodal_2025_b200.metrics.topk import top_k_score
    n_im, cap, d = 500, 5, 128
    g = torch.Generator().manual_seed(2)
    im = torch.nn.functional.normalize(torch.randn(n_im, d, generator=g), dim=-1)
    tx = torch.nn.functional.normalize(
        im.repeat_interleave(cap, 0) + 1.2 * torch.randn(n_im * cap, d, generator=g), dim=-1)
    t2i = torch.arange(n_im).repeat_interleave(cap)
    i2t = torch.arange(n_im * cap).view(n_im, cap)
    imb, txb = im.bfloat16().cuda(), tx.bfloat16().cuda()
    ks = [1, 5, 10, 50]
    got = recall_at_k_from_encodings(imb, txb, t2i, i2t, ks, fused=True)
    want = MO.recall_at_k(imb.float().cpu(), txb.float().cpu(), t2i, i2t, ks)
    for a, b in zip(got[0] + got[1], list(want[0]) + list(want[1])):
        assert abs(a - b) <= 4e-3
    pairs = [([txb[i * cap + c] for c in range(cap)], imb[i]) for i in range(n_im)]
    for k in (1, 5, 10):
        acc_f, info_f = top_k_score(pairs, k, fused=True)
        acc_p, info_p = top_k_score(pairs, k, fused=False)
        assert info_f["total"] == info_p["total"] == n_im
        assert abs(info_f["correct"] - info_p["correct"]) <= 2

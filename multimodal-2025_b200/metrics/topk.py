"""Drop-in for the reference `metrics/topk.py` (top_k_similarities :30-51, top_k_score :54-82,
TopKMetric :7-27).  The index selection runs in `mm25_topk_rows` (exact radix select, bit-exact
indices on tie-free fp32 scores); `top_k_score` scores ALL images against all texts at once
instead of the reference's Python loop over images with an `.item()` per index."""
from typing import Any, List, Tuple

import torch

from .. import ops


class TopKMetric:
    """Metric to compute the top-k accuracy based on embeddings (reference :7-27)."""

    def __init__(self, k=5):
        self.k = k

    def compute(self, embeddings: List[Tuple[List[torch.Tensor], torch.Tensor]]
                ) -> Tuple[float, Any]:
        return top_k_score(embeddings, self.k)


def top_k_similarities(embeddings, query_embedding, k=5):
    """List of (index, similarity) of the k texts most similar to the query (reference :30-51)."""
    similarities = torch.nn.functional.cosine_similarity(
        embeddings, query_embedding.unsqueeze(0), dim=1)
    idx, val = ops.topk_rows(similarities, k)
    return list(zip(idx[0].tolist(), val[0].tolist()))


# score slab held at once by top_k_indices_batched (fp32 bytes); 180 GB of HBM allow large slabs
MAX_SCORE_BYTES = 16 << 30


def top_k_indices_batched(text_emb: torch.Tensor, image_emb: torch.Tensor, k: int,
                          max_score_bytes: int = None):
    """indices [Q, k] of the top-k texts for every image (cosine similarity, eps 1e-8).  The
    [Q, N_t] scores are produced in query slabs of at most `max_score_bytes` (rows are
    independent, so the split does not change the result)."""
    t = text_emb.float()
    q = image_emb.float()
    t = t / t.norm(dim=1, keepdim=True).clamp_min(1e-8)
    q = q / q.norm(dim=1, keepdim=True).clamp_min(1e-8)
    budget = MAX_SCORE_BYTES if max_score_bytes is None else int(max_score_bytes)
    rows = max(1, min(q.shape[0], budget // max(4 * t.shape[0], 1)))
    if rows >= q.shape[0]:
        idx, _ = ops.topk_rows(q @ t.T, k)
        return idx
    return torch.cat([ops.topk_rows(q[a:a + rows] @ t.T, k)[0]
                      for a in range(0, q.shape[0], rows)])


def top_k_score(embedding_pairs, k=5, fused=None):
    """Fraction of images whose correct texts intersect their top-k texts (reference :54-82).
    `fused=True` (or MM25_FUSED_RANKS=1 with bf16 CUDA embeddings): the image x text similarity GEMM and the rank
    count run in one tcgen05 kernel (`mm25_sim_rank`, no score matrix) — an image hits iff the
    best rank among its own texts is < k; cosine normalisation is done before the bf16 cast."""
    total_count = len(embedding_pairs)
    if total_count == 0:
        return 0.0, {"correct": 0, "total": 0}
    all_text = torch.stack([t for texts, _ in embedding_pairs for t in texts])
    images = torch.stack([img for _, img in embedding_pairs])
    counts = torch.tensor([len(texts) for texts, _ in embedding_pairs], device=all_text.device)
    start = torch.cumsum(counts, 0) - counts  # `sofar` of the reference loop
    from .retrieval import _use_fused, fused_ranks
    if int(counts.min()) > 0 and _use_fused(fused, images, all_text):
        t = all_text.float()
        q = images.float()
        t = (t / t.norm(dim=1, keepdim=True).clamp_min(1e-8)).to(torch.bfloat16)
        q = (q / q.norm(dim=1, keepdim=True).clamp_min(1e-8)).to(torch.bfloat16)
        best = None
        for c in range(int(counts.max())):
            tgt = start + torch.clamp(counts - 1, max=c)
            r = fused_ranks(q, t, tgt)
            best = r if best is None else torch.minimum(best, r)
        correct_count = int((best < k).sum().item())
        return correct_count / total_count, {"correct": correct_count, "total": total_count}
    idx = top_k_indices_batched(all_text, images, k)
    hit = ((idx >= start[:, None]) & (idx < (start + counts)[:, None])).any(dim=1)
    correct_count = int(hit.sum().item())
    return correct_count / total_count, {"correct": correct_count, "total": total_count}

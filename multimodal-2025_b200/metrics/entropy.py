"""Drop-in for the reference `metrics/entropy.py` (EntropyMetric :10-34,
average_kth_neighbor_angle :37-80): the average angle between each embedding and its k-th nearest
neighbour.

Same batching and the same expressions as the reference (normalise, `mm` of a batch against all
embeddings, self-similarity -> -inf, k-th largest similarity, `acos(clamp)`, mean), so the
scores are the ones the reference computes on the same device.  The k-th largest value per row
comes from `mm25_topk_rows` (exact radix select — a value, so tie order cannot matter), the
self-mask is one indexed store instead of the reference's Python loop over the batch rows
(:66-67), and the angles stay on the device until the final `.item()`.
"""
from typing import List, Tuple

import torch
from torch import Tensor

from .. import ops as _ops  # tests swap this for the dense CPU stand-in (tests/cpu_backend.py)


class EntropyMetric:
    """Entropy proxy of a set of (text_embedding[], image_embedding) pairs (reference :10-34)."""

    def __init__(self, k=5):
        self.k = k

    def compute(self, embeddings: List[Tuple[List[Tensor], Tensor]]) -> float:
        all_embeddings = torch.cat(
            [torch.stack(text_embs + [img_emb]) for text_embs, img_emb in embeddings], dim=0)
        return average_kth_neighbor_angle(all_embeddings, self.k)


def kth_neighbor_angles(embeddings: Tensor, k=5, batch_size=1000) -> Tensor:
    """[N] angles (radians) to the k-th nearest neighbour, self excluded (reference :51-77)."""
    n = embeddings.size(0)
    if not 1 <= k <= n:
        raise RuntimeError("selected index k out of range")  # what torch.topk raises
    embeddings_norm = torch.nn.functional.normalize(embeddings, p=2, dim=1)
    out = torch.empty((n,), dtype=torch.float32, device=embeddings.device)
    for i in range(0, n, batch_size):
        batch_end = min(i + batch_size, n)
        similarities = torch.mm(embeddings_norm[i:batch_end], embeddings_norm.t()).float()
        r = torch.arange(batch_end - i, device=embeddings.device)
        similarities[r, r + i] = -float("inf")
        _, kth = _ops.topk_rows(similarities, k)
        out[i:batch_end] = torch.acos(torch.clamp(kth[:, -1], -1.0, 1.0))
    return out


def average_kth_neighbor_angle(embeddings: Tensor, k=5, batch_size=1000) -> float:
    """Average angle to the k-th nearest neighbour in radians (reference :37-80)."""
    return kth_neighbor_angles(embeddings, k, batch_size).mean().item()

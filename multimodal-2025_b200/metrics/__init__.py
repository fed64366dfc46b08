"""Mirror of the reference's retrieval-side metrics on this path: `metrics/topk.py`,
`metrics/entropy.py`, and the rank metrics of the AlignCLIP trainer (`main/train.py`
get_clip_metrics, `main/retrieval.py` recall_at_k)."""
from .topk import TopKMetric, top_k_score, top_k_similarities  # noqa: F401
from .entropy import EntropyMetric, average_kth_neighbor_angle, kth_neighbor_angles  # noqa: F401
from .retrieval import (get_clip_metrics, paired_ranks, recall_at_k,  # noqa: F401
                        recall_at_k_from_encodings, retrieval_ranks)

"""Mirror of the reference `metrics/` package: only the top-k retrieval metric is on this path."""
from .topk import TopKMetric, top_k_score, top_k_similarities  # noqa: F401

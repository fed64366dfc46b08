"""CPU oracle for the retrieval metrics — TEST INFRASTRUCTURE ONLY (imported by tests/; never by
the product path).

Restates, with the sort the reference uses, `AlignCLIP/main/train.py:338-356 get_clip_metrics`,
`AlignCLIP/main/retrieval.py:55-102 recall_at_k` (after `encode_dataset`) and
`metrics/entropy.py:37-80 average_kth_neighbor_angle`.  Pinned against the reference itself:
tests/golden/metrics.npz was produced by tests/golden/make_golden_metrics.py executing those
reference functions from /root/reference; tests/test_oracle_metrics.py checks this file against
it.  Sorts here are STABLE so that tied scores have a defined answer (lowest index first) — on
tie-free scores, the only case where the reference's answer is defined, the two agree.
"""
import numpy as np
import torch


def ranks_by_sort(scores, targets):
    """position of targets[q] in the stable descending sort of row q (train.py:349-350)."""
    ranking = torch.argsort(scores, dim=1, descending=True, stable=True)
    return torch.where(ranking == targets.view(-1, 1))[1]


def get_clip_metrics(image_features, text_features, logit_scale):
    """train.py:338-356."""
    metrics = {}
    logits_per_image = (logit_scale * image_features @ text_features.t()).detach().cpu()  # :341
    logits_per_text = logits_per_image.t().detach().cpu()  # :342
    logits = {"image_to_text": logits_per_image, "text_to_image": logits_per_text}  # :344
    ground_truth = torch.arange(len(text_features))  # :345
    for name, logit in logits.items():
        preds = ranks_by_sort(logit.float(), ground_truth).numpy()  # :348-350
        metrics[f"{name}_mean_rank"] = preds.mean() + 1  # :351
        metrics[f"{name}_median_rank"] = np.floor(np.median(preds)) + 1  # :352
        for k in [1, 5, 10]:
            metrics[f"{name}_R@{k}"] = np.mean(preds < k)  # :353-354
    return metrics


def recall_at_k(image_encodings, text_encodings, text_to_image_map, image_to_text_map, k_vals):
    """retrieval.py:59-102 from the encodings on."""
    num_text = text_encodings.shape[0]  # :59
    num_im = image_encodings.shape[0]  # :60
    captions_per_image = image_to_text_map.shape[1]  # :61
    dist_matrix = text_encodings @ image_encodings.T  # :65
    inds = torch.argsort(dist_matrix, dim=1, descending=True, stable=True)  # :69
    text_to_image_recall = []
    for k in k_vals:  # :74-80
        topk = inds[:, :k]
        correct = torch.eq(topk, text_to_image_map.unsqueeze(-1)).any(dim=1)
        text_to_image_recall.append(correct.sum().item() / num_text)
    dist_matrix = dist_matrix.T  # :83
    inds = torch.argsort(dist_matrix, dim=1, descending=True, stable=True)  # :85
    image_to_text_recall = []
    for k in k_vals:  # :90-100
        topk = inds[:, :k]
        correct = torch.zeros((num_im,), dtype=torch.bool)
        for i in range(captions_per_image):
            contains_index = torch.eq(topk, image_to_text_map[:, i].unsqueeze(-1)).any(dim=1)
            correct = torch.logical_or(correct, contains_index)
        image_to_text_recall.append(correct.sum().item() / num_im)
    return text_to_image_recall, image_to_text_recall


def kth_neighbor_angles(embeddings, k=5, batch_size=1000):
    """entropy.py:51-77 (per-embedding angles, before the mean)."""
    n = embeddings.size(0)
    embeddings_norm = torch.nn.functional.normalize(embeddings, p=2, dim=1)  # :55
    out = []
    for i in range(0, n, batch_size):  # :60
        batch_end = min(i + batch_size, n)
        similarities = torch.mm(embeddings_norm[i:batch_end], embeddings_norm.t())  # :66
        for j in range(batch_end - i):  # :69-70
            similarities[j, i + j] = -float("inf")
        kth_similarities, _ = similarities.topk(k, dim=1)  # :73
        out.append(torch.acos(torch.clamp(kth_similarities[:, -1], -1.0, 1.0)))  # :76
    return torch.cat(out)  # :80


def average_kth_neighbor_angle(embeddings, k=5, batch_size=1000):
    """entropy.py:37-80."""
    return kth_neighbor_angles(embeddings, k, batch_size).mean().item()

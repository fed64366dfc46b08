"""CPU: the metrics oracle against golden vectors produced by executing the reference
(tests/golden/make_golden_metrics.py), and the HOST LOGIC of metrics/retrieval.py and
metrics/entropy.py (target gathering, slab split, caption passes, batching) with the dense
stand-in ops of tests/cpu_backend.py in place of the CUDA kernels."""
import numpy as np
import pytest
import torch

import cpu_backend
from conftest import load_golden
from oracle import metrics_oracle as MO

G = load_golden("metrics.npz")
GCM = sorted(k for k in G if k.startswith("gcm_"))
RAK = sorted(k for k in G if k.startswith("rak_"))
ENT = sorted(k for k in G if k.startswith("ent_n"))
METRIC_KEYS = [f"{n}_{m}" for n in ("image_to_text", "text_to_image")
               for m in ("mean_rank", "median_rank", "R@1", "R@5", "R@10")]


@pytest.fixture
def host_metrics(monkeypatch):
    """metrics modules with `_ops` swapped for the dense CPU stand-in (host logic only)."""
    from multimodal_2025_b200.metrics import entropy, retrieval
    monkeypatch.setattr(retrieval, "_ops", cpu_backend)
    monkeypatch.setattr(entropy, "_ops", cpu_backend)
    return retrieval, entropy


@pytest.mark.parametrize("tag", GCM)
def test_oracle_get_clip_metrics_vs_reference(tag):
    c = G[tag]
    img, txt = torch.from_numpy(c["img"]), torch.from_numpy(c["txt"])
    m = MO.get_clip_metrics(img, txt, float(c["logit_scale"]))
    for k in METRIC_KEYS:
        assert float(m[k]) == float(c[k]), k
    logits = float(c["logit_scale"]) * img @ txt.t()
    ar = torch.arange(img.shape[0])
    assert np.array_equal(MO.ranks_by_sort(logits, ar).numpy(), c["image_to_text_preds"])
    assert np.array_equal(MO.ranks_by_sort(logits.t(), ar).numpy(), c["text_to_image_preds"])


@pytest.mark.parametrize("tag", RAK)
def test_oracle_recall_at_k_vs_reference(tag):
    c = G[tag]
    t2i, i2t = MO.recall_at_k(torch.from_numpy(c["im"]), torch.from_numpy(c["tx"]),
                              torch.from_numpy(c["t2i_map"]), torch.from_numpy(c["i2t_map"]),
                              c["k_vals"].tolist())
    assert t2i == c["t2i_recall"].tolist() and i2t == c["i2t_recall"].tolist()


@pytest.mark.parametrize("tag", ENT)
def test_oracle_entropy_vs_reference(tag):
    c = G[tag]
    v = MO.average_kth_neighbor_angle(torch.from_numpy(c["emb"]), int(c["k"]),
                                      int(c["batch_size"]))
    assert v == float(c["value"])


@pytest.mark.parametrize("tag", GCM)
@pytest.mark.parametrize("budget", [None, 16 * 1024])
def test_host_get_clip_metrics(host_metrics, tag, budget):
    retrieval, _ = host_metrics
    c = G[tag]
    img, txt = torch.from_numpy(c["img"]), torch.from_numpy(c["txt"])
    if budget is not None:  # the two-pass slab route: several slabs, ragged last one
        assert img.shape[0] ** 2 * 4 > budget
    m = retrieval.get_clip_metrics(img, txt, float(c["logit_scale"]), max_score_bytes=budget)
    assert set(m) == set(METRIC_KEYS)
    i2t, t2i = retrieval.paired_ranks(img, txt, float(c["logit_scale"]), max_score_bytes=budget)
    if budget is None:  # identical GEMM -> identical scores -> the reference's own numbers
        for k in METRIC_KEYS:
            assert float(m[k]) == float(c[k]), k
        assert np.array_equal(i2t.numpy(), c["image_to_text_preds"])
        assert np.array_equal(t2i.numpy(), c["text_to_image_preds"])
    else:  # slab GEMMs may round differently from the full GEMM: exact against a sort of the
        # scores this route actually produced
        rows = max(32, min(img.shape[0], (budget // (4 * img.shape[0])) // 32 * 32))
        s = torch.cat([float(c["logit_scale"]) * img[a:a + rows] @ txt.t()
                       for a in range(0, img.shape[0], rows)])
        ar = torch.arange(img.shape[0])
        assert torch.equal(i2t.long(), MO.ranks_by_sort(s, ar))
        assert torch.equal(t2i.long(), MO.ranks_by_sort(s.t().contiguous(), ar))


@pytest.mark.parametrize("tag", RAK)
def test_host_recall_at_k(host_metrics, tag):
    retrieval, _ = host_metrics
    c = G[tag]
    args = (torch.from_numpy(c["im"]), torch.from_numpy(c["tx"]), torch.from_numpy(c["t2i_map"]),
            torch.from_numpy(c["i2t_map"]))
    t2i, i2t = retrieval.recall_at_k_from_encodings(*args, c["k_vals"].tolist())
    assert t2i == c["t2i_recall"].tolist() and i2t == c["i2t_recall"].tolist()
    got = retrieval.recall_at_k("clip", "data", "cpu", c["k_vals"].tolist(), 16,
                                encode_dataset=lambda clip, data, device, batch_size: args)
    assert got == (t2i, i2t)
    with pytest.raises(ValueError):
        retrieval.recall_at_k("clip", "data", "cpu", [1], 16)


@pytest.mark.parametrize("tag", ENT)
def test_host_entropy(host_metrics, tag):
    _, entropy = host_metrics
    c = G[tag]
    emb = torch.from_numpy(c["emb"])
    v = entropy.average_kth_neighbor_angle(emb, int(c["k"]), int(c["batch_size"]))
    assert v == float(c["value"])
    with pytest.raises(RuntimeError):
        entropy.average_kth_neighbor_angle(emb[:4], 5)


def test_host_entropy_metric_class(host_metrics):
    _, entropy = host_metrics
    pairs = [([torch.tensor([1.0, 0.0, 0.0]), torch.tensor([0.0, 1.0, 0.0])],
              torch.tensor([0.0, 0.0, 1.0])),
             ([torch.tensor([1.0, 1.0, 0.0])], torch.tensor([1.0, 0.2, 0.9]))]
    assert entropy.EntropyMetric(k=1).compute(pairs) == float(G["ent_metric"]["value_k1"])
    assert entropy.EntropyMetric(k=3).compute(pairs) == float(G["ent_metric"]["value_k3"])


def test_standin_rank_ties_and_specials():
    """the stable-order contract of mm25_rank_both on ties, signed zeros, infinities and NaN
    (the GPU test asserts the same table on the kernel)."""
    s = torch.tensor([[1.0, 3.0, 3.0, -0.0, 0.0, float("inf"), 3.0, float("-inf")],
                      [2.0, 2.0, 2.0, 2.0, 2.0, 2.0, 2.0, 2.0],
                      [float("nan"), 1.0, float("inf"), float("nan"), 0.0, -1.0, 5.0, 5.0]])
    for tcol, want in (([2, 4, 3], [2, 4, 1]), ([3, 0, 7], [6, 0, 4]), ([5, 7, 0], [0, 7, 0])):
        t = torch.tensor(tcol)
        rr = torch.zeros(3, dtype=torch.int32)
        cpu_backend.rank_both(s, 0, s.gather(1, t.view(-1, 1)).view(-1), t, None, None, rr, None)
        assert rr.tolist() == want


@pytest.mark.parametrize("n,d,budget,seed", [(67, 8, None, 0), (130, 4, 4096, 1), (33, 3, 2048, 2),
                                             (257, 6, 40000, 3)])
def test_host_paired_ranks_with_ties_any_slab_split(host_metrics, n, d, budget, seed):
    """coarsely quantised features (many exactly tied scores, some tied with the target): the
    one-shot and the two-pass slab route give the stable-sort ranks of the scores they saw."""
    retrieval, _ = host_metrics
    g = torch.Generator().manual_seed(seed)
    x = torch.round(torch.randn(n, d, generator=g))
    y = torch.round(x + 0.8 * torch.randn(n, d, generator=g))
    i2t, t2i = retrieval.paired_ranks(x, y, 2.0, max_score_bytes=budget)
    s = 2.0 * x @ y.t()                      # small integers: every GEMM split gives the same bits
    ar = torch.arange(n)
    assert (s == s.diagonal().view(-1, 1)).sum() > n      # ties with targets really occur
    assert torch.equal(i2t.long(), MO.ranks_by_sort(s, ar))
    assert torch.equal(t2i.long(), MO.ranks_by_sort(s.t().contiguous(), ar))
    m = retrieval.get_clip_metrics(x, y, 2.0, max_score_bytes=budget)
    want = MO.get_clip_metrics(x, y, 2.0)
    for k in METRIC_KEYS:
        assert float(m[k]) == float(want[k]), k


def test_host_recall_at_k_with_ties_and_uneven_maps(host_metrics):
    """tied scores and a caption map that is not the regular i*c+j layout."""
    retrieval, _ = host_metrics
    g = torch.Generator().manual_seed(9)
    num_im, cpi, d = 40, 3, 4
    im = torch.round(torch.randn(num_im, d, generator=g))
    perm = torch.randperm(num_im * cpi, generator=g)
    t2i = (torch.arange(num_im * cpi) // cpi)[perm]          # text q belongs to image t2i[q]
    tx = torch.round(im[t2i] + 0.7 * torch.randn(num_im * cpi, d, generator=g))
    i2t = torch.stack([torch.nonzero(t2i == i).view(-1) for i in range(num_im)])
    k_vals = [1, 2, 5, 17, 120]
    got = retrieval.recall_at_k_from_encodings(im, tx, t2i, i2t, k_vals)
    want = MO.recall_at_k(im, tx, t2i, i2t, k_vals)
    assert got == want

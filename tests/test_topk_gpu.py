"""GPU: top-k retrieval indices (metrics/topk.py).  Bit-exact against torch.topk on tie-free
fp32 score rows, against the reference's golden indices, and the reference's own KAT."""
import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("Q,N,k", [(4, 1000, 1), (4, 1000, 5), (3, 1000, 10), (3, 1000, 100),
                                   (2, 65536, 100), (2, 65536, 1), (5, 257, 257), (1, 3, 2),
                                   (2, 70001, 2048)])
def test_topk_indices_bit_exact_vs_torch(lib_built, Q, N, k):
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(N + k)
    # tie-free rows: a random permutation of distinct values
    base = torch.linspace(-3.0, 3.0, N)
    scores = torch.stack([base[torch.randperm(N, generator=g)] for _ in range(Q)])
    ref = torch.topk(scores, k, dim=1)
    idx, val = ops.topk_rows(scores.cuda(), k)
    assert torch.equal(idx.cpu(), ref.indices)
    assert torch.equal(val.cpu(), ref.values)


def test_topk_ties_and_special_values(lib_built):
    from multimodal_2025_b200 import ops
    s = torch.tensor([[1.0, 3.0, 3.0, -0.0, 0.0, float("inf"), 3.0, float("-inf")],
                      [2.0, 2.0, 2.0, 2.0, 2.0, 2.0, 2.0, 2.0]])
    idx, val = ops.topk_rows(s.cuda(), 4)
    assert idx[0].tolist() == [5, 1, 2, 6]          # ties: lowest index first
    assert idx[1].tolist() == [0, 1, 2, 3]
    ref = torch.topk(s, 4, dim=1)
    assert torch.equal(val.cpu(), ref.values)       # as a set of values torch agrees


def test_topk_signed_zeros_tie_like_torch(lib_built):
    """torch.topk / argsort compare VALUES: -0.0 and +0.0 tie (lowest index first here), and
    both zeros reach the top-k of these rows (metrics/topk.py:48; orthogonal or zero rows give
    exactly such scores)."""
    from multimodal_2025_b200 import ops
    s = torch.tensor([[-1.0, -0.0, 0.0, -2.0, 0.0, -0.0, -3.0, -0.5],
                      [0.0, -0.0, -0.0, 0.0, -1.0, 1.0, -0.0, 0.0]])
    idx, val = ops.topk_rows(s.cuda(), 3)
    assert idx[0].tolist() == [1, 2, 4]
    assert idx[1].tolist() == [5, 0, 1]
    order = torch.sort(s, dim=1, descending=True, stable=True).indices[:, :3]
    assert torch.equal(idx.cpu(), order)
    # the VALUES come back with their own sign bit (gathered from the row, not from the key)
    assert torch.equal(torch.signbit(val.cpu()), torch.signbit(s.gather(1, order)))
    g = torch.Generator().manual_seed(11)
    big = torch.round(torch.randn(7, 5000, generator=g) * 2) / 2
    big = torch.where(torch.rand(7, 5000, generator=g) < 0.3, -big, big)  # many -0.0
    for k in (1, 17, 700):
        idx, _ = ops.topk_rows(big.cuda(), k)
        want = torch.sort(big, dim=1, descending=True, stable=True).indices[:, :k]
        assert torch.equal(idx.cpu(), want)


def test_topk_reference_golden_and_kat(lib_built):
    from multimodal_2025_b200.metrics import top_k_score, top_k_similarities
    G = load_golden("topk.npz")
    top = G[""]
    texts = torch.from_numpy(top["texts"]).cuda()
    images = torch.from_numpy(top["images"]).cuda()
    # the reference's __main__ KAT (metrics/topk.py:87-106)
    emb1 = torch.tensor([[1.0, 0.0], [0.9, 0.1], [0.0, 1.0]]).cuda()
    kat = top_k_similarities(emb1, torch.tensor([1.0, 0.0]).cuda(), k=2)
    assert [i for i, _ in kat] == top["kat_idx"].tolist()
    assert np.allclose([s for _, s in kat], top["kat_sim"], rtol=1e-6)
    pairs = [([texts[3 * i], texts[3 * i + 1], texts[3 * i + 2]], images[i]) for i in range(100)]
    for k in (1, 5, 10, 100):
        acc, info = top_k_score(pairs, k)
        assert info["correct"] == int(G["k%d" % k]["correct"]) and info["total"] == 100
        assert acc == float(G["k%d" % k]["acc"])
        got = np.array([[i for i, _ in top_k_similarities(texts, images[q], k)]
                        for q in range(0, 100, 9)])
        assert np.array_equal(got, G["k%d" % k]["idx"][::9])


@pytest.mark.parametrize("Q,N,k", [(1, 1 << 20, 5), (1, 1000003, 100), (3, 40000, 64),
                                   (16, 300001, 10), (300, 16384, 7), (300, 16385, 2048),
                                   (2, 70001, 2048), (1, 9000, 9000 // 5)])
def test_topk_sliced_path_equals_plain_kernel_and_stable_sort(lib_built, monkeypatch, Q, N, k):
    """round 2: rows cut into shared-memory slices (one HBM pass, small-Q rows spread over the
    SMs) + merge — bit-identical to the one-CTA-per-row kernel and to the stable descending sort,
    on tie-heavy rows with signed zeros, infinities and ragged / unaligned lengths."""
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(N + k + Q)
    s = torch.round(torch.randn(Q, N, generator=g) * 8) / 8      # ~60 distinct values: heavy ties
    s = torch.where(torch.rand(Q, N, generator=g) < 0.2, -s, s)   # signed zeros
    s[:, ::7919] = float("inf")
    s[:, 5::10007] = float("-inf")
    sc = s.cuda()
    idx, val = ops.topk_rows(sc, k)
    monkeypatch.setenv("MM25_TOPK_PLAIN", "1")
    idx_p, val_p = ops.topk_rows(sc, k)
    monkeypatch.delenv("MM25_TOPK_PLAIN")
    assert torch.equal(idx, idx_p) and torch.equal(val, val_p)
    want = torch.sort(s, dim=1, descending=True, stable=True).indices[:, :k]
    assert torch.equal(idx.cpu(), want)
    # an unaligned view (row pointer not 16-byte aligned): the scalar load path
    v = sc[:, 1:]
    idx_v, _ = ops.topk_rows(v, min(k, N - 1))
    want_v = torch.sort(s[:, 1:], dim=1, descending=True, stable=True).indices[:, :min(k, N - 1)]
    assert torch.equal(idx_v.cpu(), want_v)


@pytest.mark.parametrize("Q,N,k", [(300, 40000, 1), (300, 40000, 100), (296, 65536, 2048),
                                   (310, 20001, 37)])
def test_topk_compaction_path_many_long_rows(lib_built, monkeypatch, Q, N, k):
    """round 2: many long rows take the two-pass kernel (12-bit histogram, bin compacted into
    shared memory); rows whose bin overflows (heavy ties) are repeated by the five-pass kernel.
    Mixed here: even rows tie-free (compaction), odd rows tie-heavy with signed zeros / inf
    (hand-back) — all bit-identical to the plain kernel and to the stable descending sort."""
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(N + k)
    base = torch.linspace(-4.0, 4.0, N)
    s = torch.stack([base[torch.randperm(N, generator=g)] for _ in range(Q)])
    t = torch.round(torch.randn(Q, N, generator=g) * 4) / 4
    t = torch.where(torch.rand(Q, N, generator=g) < 0.2, -t, t)
    t[:, ::5003] = float("inf")
    s[1::2] = t[1::2]
    s[0, :50] = 3.999                      # a run of equal keys inside a compacted bin
    sc = s.cuda()
    idx, val = ops.topk_rows(sc, k)
    monkeypatch.setenv("MM25_TOPK_PLAIN", "1")
    idx_p, val_p = ops.topk_rows(sc, k)
    monkeypatch.delenv("MM25_TOPK_PLAIN")
    assert torch.equal(idx, idx_p) and torch.equal(val, val_p)
    want = torch.sort(s, dim=1, descending=True, stable=True).indices[:, :k]
    assert torch.equal(idx.cpu(), want)

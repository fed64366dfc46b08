"""GPU parity of the CLIP loss (through the C-ABI) against the reference's golden vectors and
the CPU oracle.  Tolerances are BASELINE.json's: fp32 rel 1e-5 loss / 1e-4 grads; bf16 rel 2e-3
loss / 2e-2 grads."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import losses_oracle as O

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


def run_ours(mod_fn, tensors, dtype=torch.float32):
    leaves = [t.detach().to("cuda").to(dtype if t.dim() > 0 else torch.float32)
              .requires_grad_(True) for t in tensors]
    out = mod_fn(*leaves)
    out.sum().backward()
    return out.detach().float().cpu(), [l.grad.float().cpu() for l in leaves]


@pytest.mark.parametrize("case", ["n256_d512", "n256_d512_clamp", "n192_d96_eps0", "n70_d40_raw",
                                  "n96_d64_nonorm"])
def test_clip_fp32_vs_reference_golden(lib_built, case):
    from multimodal_2025_b200.losses import ClipLoss
    g = load_golden("clip_forward.npz")[case]
    norm = case != "n96_d64_nonorm"
    m = ClipLoss(float(g["eps"])).cuda()
    loss, (gx, gy, gl) = run_ours(lambda a, b, c: m(a, b, c, normalize=norm),
                                  [T(g["x"]), T(g["y"]), torch.tensor(float(g["ls"]))])
    assert rel_err(loss, g["loss"]) < 1e-5
    assert rel_err(gx, g["gx"]) < 1e-4
    assert rel_err(gy, g["gy"]) < 1e-4
    assert abs(float(gl) - float(g["gl"])) <= 1e-4 * max(abs(float(g["gl"])), 1e-3)


@pytest.mark.parametrize("n,d,seed,ls", [(64, 64, 0, 2.0), (257, 100, 1, 3.0), (1000, 512, 2, 4.6052),
                                         (33, 7, 3, 1.0), (1, 16, 4, 2.0)])
def test_clip_fp32_vs_oracle_shapes(lib_built, n, d, seed, ls):
    """ragged / tiny / single-row shapes and the clamp boundary."""
    from multimodal_2025_b200.losses import ClipLoss
    x, y = O.synth_embeddings(n, d, seed, correlated=True)
    lsc = torch.tensor(ls)
    ref, (rx, ry, rl) = O.loss_and_grads(O.clip_loss, [x, y, lsc], 0.1, True)
    m = ClipLoss(0.1).cuda()
    loss, (gx, gy, gl) = run_ours(lambda a, b, c: m(a, b, c), [x, y, lsc])
    assert abs(float(loss) - float(ref)) <= 1e-5 * max(abs(float(ref)), 1e-3)
    if n > 1:
        assert rel_err(gx, rx) < 1e-4 and rel_err(gy, ry) < 1e-4
        assert abs(float(gl) - float(rl)) <= 1e-4 * max(abs(float(rl)), 1e-3)


@pytest.mark.parametrize("key", ["eps0.1_ls2.6592", "eps0.1_ls4.2", "eps0.0_ls2.6592",
                                 "eps0.0_ls4.2"])
def test_clip_sampled_fp32_vs_reference_golden(lib_built, key):
    from multimodal_2025_b200.losses import ClipLoss
    G = load_golden("clip_sampled.npz")
    g, top = G[key], G[""]
    eps = float(key.split("_")[0][3:])
    ls = float(key.split("ls")[1])
    w = T(top["w"]).cuda()
    m = ClipLoss(eps).cuda()
    loss, (gx, gy, gl) = run_ours(lambda a, b, c: m.forward_sampled(a, b, c) * w,
                                  [T(top["x"]), T(top["y"]), torch.tensor(ls)])
    assert rel_err(loss / w.cpu(), g["loss"]) < 1e-5
    assert rel_err(gx, g["gx"]) < 1e-4 and rel_err(gy, g["gy"]) < 1e-4
    assert abs(float(gl) - float(g["gl"])) <= 1e-4 * max(abs(float(g["gl"])), 1e-3)


@pytest.mark.parametrize("impl", [1, 0])
@pytest.mark.parametrize("n,d,ls", [(256, 512, 2.6592), (512, 128, 4.6052), (384, 64, 3.5),
                                    (1024, 256, 3.0), (2048, 512, 4.6052)])
def test_clip_bf16_vs_oracle(lib_built, impl, n, d, ls):
    """bf16 inputs: SIMT (impl=1) and auto (tcgen05 when eligible) against the fp32 oracle fed
    the same bf16-rounded values."""
    from multimodal_2025_b200.losses import ClipLoss
    x, y = O.synth_embeddings(n, d, 5, correlated=True, dtype=torch.bfloat16)
    lsc = torch.tensor(ls)
    ref, (rx, ry, rl) = O.loss_and_grads(O.clip_loss, [x.float(), y.float(), lsc], 0.1, True)
    m = ClipLoss(0.1).cuda()
    m.impl = impl
    loss, (gx, gy, gl) = run_ours(lambda a, b, c: m(a, b, c), [x, y, lsc], torch.bfloat16)
    assert rel_err(loss, ref) < 2e-3
    assert rel_err(gx, rx) < 2e-2 and rel_err(gy, ry) < 2e-2
    assert abs(float(gl) - float(rl)) <= 2e-2 * max(abs(float(rl)), 1e-3)


@pytest.mark.parametrize("n,d,ls", [(256, 512, 2.6592), (1024, 256, 3.0), (2048, 512, 4.6052)])
def test_clip_bf16_kept_e_vs_oracle(lib_built, monkeypatch, n, d, ls):
    """opt-in kept-E backward (MM25_KEEP_E_GB > 0; the last case exceeds the static-shift range,
    so the flag stays down and the backward recomputes) against the same oracle and contract."""
    from multimodal_2025_b200.losses import ClipLoss
    monkeypatch.setenv("MM25_KEEP_E_GB", "1")
    x, y = O.synth_embeddings(n, d, 5, correlated=True, dtype=torch.bfloat16)
    lsc = torch.tensor(ls)
    ref, (rx, ry, rl) = O.loss_and_grads(O.clip_loss, [x.float(), y.float(), lsc], 0.1, True)
    m = ClipLoss(0.1).cuda()
    loss, (gx, gy, gl) = run_ours(lambda a, b, c: m(a, b, c), [x, y, lsc], torch.bfloat16)
    assert rel_err(loss, ref) < 2e-3
    assert rel_err(gx, rx) < 2e-2 and rel_err(gy, ry) < 2e-2
    assert abs(float(gl) - float(rl)) <= 2e-2 * max(abs(float(rl)), 1e-3)


def test_clip_linearity_in_upstream_grad(lib_built):
    """size-independent property: grads scale linearly with the upstream gradient."""
    from multimodal_2025_b200.losses import ClipLoss
    x, y = O.synth_embeddings(300, 96, 9)
    m = ClipLoss(0.1).cuda()
    _, (g1, _, _) = run_ours(lambda a, b, c: m(a, b, c), [x, y, torch.tensor(3.0)])
    _, (g3, _, _) = run_ours(lambda a, b, c: 3.0 * m(a, b, c), [x, y, torch.tensor(3.0)])
    assert rel_err(g3, 3.0 * g1) < 1e-6


def test_clip_bf16_tc_multi_panel_and_sampled(lib_built):
    """tensor-core backward with several row panels (G workspace capped) and a batched
    forward_sampled call; compared with the SIMT path on the same inputs."""
    from multimodal_2025_b200 import _lib
    from multimodal_2025_b200.losses import ClipLoss
    L = _lib.lib()
    g = torch.Generator().manual_seed(3)
    xs = torch.randn(3, 512, 256, generator=g)
    ys = xs + 2.0 * torch.randn(3, 512, 256, generator=g)
    w = torch.tensor([1.0, 0.5, 2.0]).cuda()
    outs = []
    # SIMT reference; per-entry launches with 2 row panels; the batched launch (all 3 sample
    # slices in one gradient-panel kernel + one batched grouped GEMM, per-entry upstream weights)
    for impl, cap in ((1, 0), (0, 256 * 512 * 2), (0, 0)):
        L.mm25_debug_set_g_panel_bytes(cap)
        try:
            m = ClipLoss(0.1).cuda()
            m.impl = impl
            outs.append(run_ours(lambda a, b, c: m.forward_sampled(a, b, c) * w,
                                 [xs, ys, torch.tensor(3.0)], torch.bfloat16))
        finally:
            L.mm25_debug_set_g_panel_bytes(0)
    l0, (gx0, gy0, gl0) = outs[0]
    for l1, (gx1, gy1, gl1) in outs[1:]:
        assert rel_err(l1, l0) < 2e-3
        assert rel_err(gx1, gx0) < 2e-2 and rel_err(gy1, gy0) < 2e-2
        assert abs(float(gl1) - float(gl0)) <= 2e-2 * abs(float(gl0))
    # the two tensor-core variants compute the same thing in a different launch structure
    assert rel_err(outs[2][1][0], outs[1][1][0]) < 1e-3


def test_clip_fp32_inputs_bf16_compute(lib_built):
    """compute_dtype=bf16: fp32 inputs on the tcgen05 kernels (bf16 contract), forward and
    forward_sampled; also picked up under torch.autocast(bf16)."""
    from multimodal_2025_b200.losses import ClipLoss
    x, y = O.synth_embeddings(512, 256, 21, correlated=True)
    lsc = torch.tensor(2.6592)
    ref, (rx, ry, rl) = O.loss_and_grads(O.clip_loss, [x, y, lsc], 0.1, True)
    m = ClipLoss(0.1).cuda()
    m.compute_dtype = torch.bfloat16
    loss, (gx, gy, gl) = run_ours(lambda a, b, c: m(a, b, c), [x, y, lsc])
    assert rel_err(loss, ref) < 2e-3
    assert rel_err(gx, rx) < 2e-2 and rel_err(gy, ry) < 2e-2
    assert abs(float(gl) - float(rl)) <= 2e-2 * abs(float(rl))
    m2 = ClipLoss(0.1).cuda()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss2, _ = run_ours(lambda a, b, c: m2(a, b, c), [x, y, lsc])
    assert float(loss2) == float(loss)
    g = torch.Generator().manual_seed(4)
    xs = torch.randn(3, 256, 256, generator=g)
    ys = xs + 1.5 * torch.randn(3, 256, 256, generator=g)
    refs, (rxs, rys, rls) = O.loss_and_grads(lambda a, b, c: O.clip_loss_sampled(a, b, c, 0.1),
                                             [xs, ys, lsc])
    ls_, (gxs, gys, gls) = run_ours(lambda a, b, c: m.forward_sampled(a, b, c), [xs, ys, lsc])
    assert rel_err(ls_, refs) < 2e-3
    assert rel_err(gxs, rxs) < 2e-2 and rel_err(gys, rys) < 2e-2


def test_graphed_step_and_input_pipeline(lib_built):
    """graphs.GraphedLossStep: replay equals the eager step on new inputs, and the double-buffered
    host->device pipeline (prefetch / run_prefetched) returns the batches in order."""
    from multimodal_2025_b200.graphs import GraphedLossStep
    from multimodal_2025_b200.losses import ClipLoss
    n, d = 512, 256
    crit = ClipLoss(0.1).cuda()
    batches = [O.synth_embeddings(n, d, 20 + i, correlated=True, dtype=torch.bfloat16)
               for i in range(4)]
    ls = torch.tensor(2.6592, device="cuda")
    x0, y0 = batches[0]
    step = GraphedLossStep(crit, x0.cuda(), y0.cuda(), ls)
    eager = []
    for x, y in batches:
        xc, yc = x.cuda().requires_grad_(True), y.cuda().requires_grad_(True)
        lc = ls.clone().requires_grad_(True)
        loss = crit(xc, yc, lc)
        loss.backward()
        eager.append((loss.detach().clone(), xc.grad.clone(), yc.grad.clone(), lc.grad.clone()))
    for (x, y), ref in zip(batches, eager):          # plain replay
        got = step(x.cuda(), y.cuda())
        torch.cuda.synchronize()
        for g, r in zip(got, ref):
            assert torch.equal(g, r)
    host = [(x.pin_memory(), y.pin_memory()) for x, y in batches]
    step.prefetch(*host[0])
    for t, ref in enumerate(eager):                  # pipelined replay
        if t + 1 < len(host):
            step.prefetch(*host[t + 1])
        got = [g.clone() for g in step.run_prefetched()]
        torch.cuda.synchronize()
        for g, r in zip(got, ref):
            assert torch.equal(g, r)
    with pytest.raises(RuntimeError):
        step.run_prefetched()

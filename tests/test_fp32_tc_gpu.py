"""GPU: the fp32-grade CLIP path on the tensor cores (two-term bf16 splits swept over K = 3 d,
csrc/sim_tc_bwd.cu launch_bwd_split; reference losses/clipLoss.py:106 computes fp32 logits).
Contract of the fp32 path: loss rel 1e-5, gradients rel 1e-4 against the CPU oracle — the same
bar the fp32 FFMA kernels are held to, here at 10-30x their speed."""
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import losses_oracle as O

pytestmark = pytest.mark.gpu


def _ours(fn, tensors):
    leaves = [t.detach().clone().cuda().requires_grad_(True) for t in tensors]
    out = fn(*leaves)
    out.sum().backward()
    torch.cuda.synchronize()
    return out.detach().cpu(), [l.grad.cpu() for l in leaves]


def test_split_kernel_layout_and_precision(lib_built):
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 70, 96, generator=g).cuda()
    for role in (0, 1):
        o4, hi, lo = ops.split_bf16x2(x, role)
        assert o4.shape == (2, 70, 3 * 96) and hi.dtype == torch.bfloat16
        assert torch.equal(hi, x.bfloat16())
        assert torch.equal(lo, (x - hi.float()).bfloat16())
        rec = hi.float() + lo.float()
        assert float(((rec - x).abs() / x.abs().clamp_min(1e-20)).max()) < 2.0 ** -15
        parts = o4.view(2, 70, 3, 96)
        want = (hi, hi, lo) if role == 0 else (hi, lo, hi)   # the lo x lo product is dropped
        for k in range(3):
            assert torch.equal(parts[:, :, k], want[k])


@pytest.mark.parametrize("n,d,ls,eps", [(256, 512, 2.6592, 0.1), (512, 256, 3.5, 0.0),
                                        (1024, 512, 4.6052, 0.1), (768, 1024, 2.0, 0.1)])
def test_clip_fp32_split_path_vs_oracle(lib_built, monkeypatch, n, d, ls, eps):
    """fp32 inputs of an eligible shape run the split tensor-core path (checked through the
    library's last-path probe) and meet the fp32 contract; MM25_FP32_TC=0 (FFMA kernels) agrees."""
    from multimodal_2025_b200 import _lib
    from multimodal_2025_b200.losses import ClipLoss
    x, y = O.synth_embeddings(n, d, 5, correlated=True)
    lsc = torch.tensor(ls)
    ref, (rx, ry, rl) = O.loss_and_grads(O.clip_loss, [x, y, lsc], eps, True)
    m = ClipLoss(eps).cuda()
    loss, (gx, gy, gl) = _ours(lambda a, b, c: m(a, b, c), [x, y, lsc])
    assert _lib.lib().mm25_debug_last_sim_path() == 1          # tensor-core kernels ran
    assert rel_err(loss, ref) < 1e-5
    assert rel_err(gx, rx) < 1e-4 and rel_err(gy, ry) < 1e-4
    assert abs(float(gl) - float(rl)) <= 1e-4 * max(abs(float(rl)), 1e-3)
    monkeypatch.setenv("MM25_FP32_TC", "0")
    loss2, (gx2, gy2, gl2) = _ours(lambda a, b, c: m(a, b, c), [x, y, lsc])
    assert rel_err(loss2, ref) < 1e-5 and rel_err(gx2, rx) < 1e-4


@pytest.mark.parametrize("case", ["n256_d512", "n256_d512_clamp"])
def test_clip_fp32_split_reference_golden(lib_built, case):
    """the reference-produced fp32 goldens of an eligible shape go through the split path."""
    import numpy as np
    from multimodal_2025_b200 import _lib
    from multimodal_2025_b200.losses import ClipLoss
    g = load_golden("clip_forward.npz")[case]
    x, y = torch.from_numpy(np.asarray(g["x"])), torch.from_numpy(np.asarray(g["y"]))
    m = ClipLoss(float(g["eps"])).cuda()
    loss, (gx, gy, gl) = _ours(lambda a, b, c: m(a, b, c), [x, y, torch.tensor(float(g["ls"]))])
    assert _lib.lib().mm25_debug_last_sim_path() == 1
    assert rel_err(loss, g["loss"]) < 1e-5
    assert rel_err(gx, g["gx"]) < 1e-4 and rel_err(gy, g["gy"]) < 1e-4
    assert abs(float(gl) - float(g["gl"])) <= 1e-4 * max(abs(float(g["gl"])), 1e-3)


def test_forward_sampled_fp32_split_vs_oracle(lib_built):
    from multimodal_2025_b200.losses import ClipLoss
    S, B, d = 3, 512, 256
    g = torch.Generator().manual_seed(4)
    x = torch.randn(S, B, d, generator=g)
    y = x + 0.6 * torch.randn(S, B, d, generator=g)
    w = torch.rand(S, generator=g) + 0.5
    ls = torch.tensor(3.0)

    def ref_fn(a, b, c):
        return (O.clip_loss_sampled(a, b, c, 0.1) * w).sum()
    ref, (rx, ry, rl) = O.loss_and_grads(ref_fn, [x, y, ls])
    m = ClipLoss(0.1).cuda()
    wc = w.cuda()
    loss, (gx, gy, gl) = _ours(lambda a, b, c: (m.forward_sampled(a, b, c) * wc).sum(), [x, y, ls])
    assert rel_err(loss, ref) < 1e-5
    assert rel_err(gx, rx) < 1e-4 and rel_err(gy, ry) < 1e-4
    assert abs(float(gl) - float(rl)) <= 1e-4 * max(abs(float(rl)), 1e-3)


def test_clip_fp32_split_full_size_slab(lib_built):
    """cfg4-sized slice (B=8192, d=512) in fp32: the loss against a blockwise fp64 evaluation of the
    reference formula on the GPU, and a 256-row gradient slab against the closed form."""
    from multimodal_2025_b200.losses import ClipLoss
    N, d, LS, EPS = 8192, 512, 2.6592, 0.1
    x, y = O.synth_embeddings(N, d, 77, correlated=True)
    xc = x.cuda().requires_grad_(True)
    yc = y.cuda().requires_grad_(True)
    ls = torch.tensor(LS, device="cuda", requires_grad=True)
    loss = ClipLoss(EPS).cuda()(xc, yc, ls)
    loss.backward()
    xd, yd = x.cuda().double(), y.cuda().double()
    s = float(torch.tensor(LS).exp())
    z = s * (xd @ yd.T)
    r = torch.logsumexp(z, dim=1)
    c = torch.logsumexp(z, dim=0)
    ref = ((r.sum() + c.sum()) / (2 * N) - (1 - EPS) / N * z.diagonal().sum()
           - EPS / N ** 2 * z.sum())
    assert abs(float(loss) - float(ref)) < 1e-5 * abs(float(ref))
    rows = slice(3000, 3256)
    G = (0.5 / N) * ((z[rows] - r[rows, None]).exp() + (z[rows] - c[None, :]).exp()) - EPS / N ** 2
    idx = torch.arange(3000, 3256, device="cuda")
    G[torch.arange(256, device="cuda"), idx] -= (1 - EPS) / N
    gx_hat = s * (G @ yd)
    xr = xd[rows]                       # already unit norm: d/dx of normalise at |x| = 1
    want = gx_hat - xr * (xr * gx_hat).sum(-1, keepdim=True)
    assert rel_err(xc.grad[rows].double().cpu(), want.cpu()) < 1e-4


@pytest.mark.parametrize("n,d", [(512, 256), (1024, 512), (768, 1024)])
def test_cloob_fp32_split_path_vs_oracle(lib_built, monkeypatch, n, d):
    """CLOOB with fp32 inputs of an eligible shape: the four Hopfield retrievals (forward P K and
    the dual-accumulator backward) and both InfoLOOB terms run the split tensor-core kernels and
    meet the fp32 contract; MM25_FP32_TC=0 (FFMA kernels) agrees."""
    from multimodal_2025_b200 import _lib
    from multimodal_2025_b200.losses import CLOOBLoss
    x, y = O.synth_embeddings(n, d, 8, correlated=True)
    ref, (rx, ry) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0), [x, y])
    crit = CLOOBLoss(30.0, 15.0, "cuda")
    loss, (gx, gy) = _ours(lambda a, b: crit(a, b), [x, y])
    assert _lib.lib().mm25_debug_last_sim_path() == 1
    assert rel_err(loss, ref) < 1e-5
    assert rel_err(gx, rx) < 1e-4 and rel_err(gy, ry) < 1e-4
    monkeypatch.setenv("MM25_FP32_TC", "0")
    loss2, (gx2, gy2) = _ours(lambda a, b: crit(a, b), [x, y])
    assert _lib.lib().mm25_debug_last_sim_path() == 0
    assert rel_err(loss2, ref) < 1e-5 and rel_err(gx2, rx) < 1e-4


def test_hopfield_fp32_split_pieces_vs_dense(lib_built):
    """hopfield(Q, K) and its gradients on a rectangular problem against dense fp64 autograd."""
    from multimodal_2025_b200.losses import hopfield
    g = torch.Generator().manual_seed(6)
    q = torch.nn.functional.normalize(torch.randn(256, 256, generator=g), dim=-1)
    k = torch.nn.functional.normalize(torch.randn(512, 256, generator=g), dim=-1)
    w = torch.randn(256, 256, generator=g)

    def dense(a, b):
        p = torch.softmax(15.0 * a @ b.T, dim=1) @ b
        return torch.nn.functional.normalize(p, dim=-1)
    a, b = q.double().requires_grad_(True), k.double().requires_grad_(True)
    (dense(a, b) * w.double()).sum().backward()
    qc, kc = q.cuda().requires_grad_(True), k.cuda().requires_grad_(True)
    u = hopfield(qc, kc, 15.0).T
    (u * w.cuda()).sum().backward()
    assert rel_err(u.detach().cpu(), dense(q.double(), k.double())) < 1e-5
    assert rel_err(qc.grad.cpu(), a.grad) < 1e-4 and rel_err(kc.grad.cpu(), b.grad) < 1e-4

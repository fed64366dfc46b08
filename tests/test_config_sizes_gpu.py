"""GPU parity at the sizes BASELINE.json names for the configs other than cfg2 (which has
tests/test_full_size_gpu.py): cfg3 CLOOB N=4096 d=512 against the CPU oracle run whole; cfg4
forward_sampled S=20, B=8192 d=512 against the CPU oracle on two of the twenty (independent)
sample slices; the cfg5 per-GPU block (M=8192 rows of N=65536, d=1024, d_sem=768, MODULATED
logits, positives at a column offset) against a direct fp32 evaluation of a slab of rows / columns
of the reference formula (AlignCLIP/align_clip/loss.py:229-242).
Tolerances: fp32 path rel 1e-5 loss / 1e-4 gradients; bf16 path 2e-3 / 2e-2 (BASELINE.json)."""
import pytest
import torch

from conftest import rel_err
from oracle import losses_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cloob_ref():
    x, y = O.synth_embeddings(4096, 512, 21, correlated=True)
    ref, (rx, ry) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0), [x, y])
    return x, y, ref, rx, ry


@pytest.mark.parametrize("dtype,tl,tg", [(torch.float32, 1e-5, 1e-4), (torch.bfloat16, 2e-3, 2e-2)])
def test_cfg3_cloob_n4096_vs_oracle(lib_built, cloob_ref, dtype, tl, tg):
    from multimodal_2025_b200.losses import CLOOBLoss
    x, y, ref, rx, ry = cloob_ref
    if dtype == torch.bfloat16:  # the oracle sees the same bf16-rounded inputs
        xb, yb = x.bfloat16(), y.bfloat16()
        ref, (rx, ry) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0),
                                         [xb.float(), yb.float()])
        x, y = xb, yb
    xc = x.cuda().requires_grad_(True)
    yc = y.cuda().requires_grad_(True)
    loss = CLOOBLoss(30.0, 15.0, "cuda")(xc, yc)
    loss.backward()
    assert rel_err(loss.detach().float().cpu(), ref) < tl
    assert rel_err(xc.grad.float().cpu(), rx) < tg and rel_err(yc.grad.float().cpu(), ry) < tg


def test_cfg4_forward_sampled_s20_b8192_slices_vs_oracle(lib_built):
    """S=20, B=8192, d=512 through the bf16 tensor-core kernels (one batched launch set); the
    slices are independent, so slices 0 and 19 are checked against the CPU oracle fed the same
    bf16-rounded samples: loss[s] and the gradient of sum_s w_s loss[s] w.r.t. those slices."""
    from multimodal_2025_b200.losses import ClipLoss
    S, B, d = 20, 8192, 512
    g = torch.Generator().manual_seed(4)
    x = torch.nn.functional.normalize(torch.randn(S, B, d, generator=g), dim=-1).bfloat16()
    y = torch.nn.functional.normalize(
        x.float() + 0.6 * torch.randn(S, B, d, generator=g), dim=-1).bfloat16()
    w = torch.rand(S, generator=g) + 0.5
    ls = torch.tensor(2.6592)
    m = ClipLoss(0.1).cuda()
    xc = x.cuda().requires_grad_(True)
    yc = y.cuda().requires_grad_(True)
    lc = ls.cuda().requires_grad_(True)
    loss = m.forward_sampled(xc, yc, lc)
    (loss * w.cuda()).sum().backward()
    torch.cuda.synchronize()
    assert loss.shape == (S,)
    gl = 0.0
    for s in (0, 19):
        xs = x[s:s + 1].float().requires_grad_(True)
        ys = y[s:s + 1].float().requires_grad_(True)
        l = ls.clone().requires_grad_(True)
        ref = O.clip_loss_sampled(xs, ys, l, 0.1)
        (ref * w[s]).sum().backward()
        assert rel_err(loss[s].detach().cpu(), ref.detach()[0]) < 2e-3
        assert rel_err(xc.grad[s].float().cpu(), xs.grad[0]) < 2e-2
        assert rel_err(yc.grad[s].float().cpu(), ys.grad[0]) < 2e-2
        gl += float(l.grad)
    # d/dlogit_scale sums over all 20 slices; the two checked ones bound it from one side only,
    # so check the mean-per-slice magnitude instead (slices are identically distributed)
    per_slice = float(lc.grad) / float(w.sum())
    assert abs(per_slice - gl / float(w[0] + w[19])) < 0.1 * abs(per_slice)


def test_cfg5_block_modulated_m8192_n65536_vs_direct(lib_built):
    """The per-GPU block of cfg5 (rank 3 of 8): rows [24576, 32768) of N=65536, d=1024, d_sem=768,
    bf16 features on the two-accumulator tcgen05 kernels.  z = s [(a_i.b_j)(1 - u_i.u_j) + delta D_i]
    (loss.py:229-242); loss = mean_i (lse_i - z_ii).  Checked on slabs against fp32 torch:
    row LSE and positives of 256 rows, dA of those rows, dB of 256 columns (needs all M rows),
    d/dD of the rows."""
    from multimodal_2025_b200 import ops
    M, N, d, ds, off = 8192, 65536, 1024, 768, 24576
    g = torch.Generator().manual_seed(9)
    dev = "cuda"
    f = torch.nn.functional.normalize(torch.randn(N, d, generator=g), dim=-1).bfloat16().to(dev)
    u = torch.nn.functional.normalize(torch.randn(N, ds, generator=g), dim=-1).to(dev)
    D = (torch.rand(M, generator=g) * 0.5 + 0.3).to(dev)
    s = 14.285
    a3, b3 = f[off:off + M].unsqueeze(0).contiguous(), f.unsqueeze(0)
    u3, v3 = u[off:off + M].unsqueeze(0).contiguous(), u.unsqueeze(0)
    sc = torch.tensor([s], device=dev)
    row_lse, _, _, diag = ops.sim_fwd(a3, b3, u3, v3, D.unsqueeze(0), sc, ops.SCALE_MUL, 0.0,
                                      ops.SIM_MODULATED, off, False, ops.IMPL_AUTO)
    gout = torch.ones(1, device=dev)
    da, db, dscale, ddiag = ops.sim_bwd(a3, b3, u3, v3, D.unsqueeze(0), sc, ops.SCALE_MUL, 0.0,
                                        ops.SIM_MODULATED, off, row_lse, None, 1.0 / M, 0.0,
                                        1.0 / M, 0.0, gout, True, True, True, ops.IMPL_AUTO)
    torch.cuda.synchronize()
    ub = u.bfloat16().float()  # the tensor-core path feeds the semantic features as bf16

    def z_rows(r0, r1):
        a = f[off + r0:off + r1].float()
        z = s * (a @ f.float().T) * (1.0 - ub[off + r0:off + r1] @ ub.T)
        idx = torch.arange(r0, r1, device=dev)
        z[idx - r0, idx + off] += s * D[r0:r1]
        return z

    r0, r1 = 1000, 1256
    z = z_rows(r0, r1)
    lse = torch.logsumexp(z, dim=1)
    idx = torch.arange(r0, r1, device=dev)
    assert rel_err(row_lse[0, r0:r1].cpu(), lse.cpu()) < 2e-3
    assert rel_err(diag[0, r0:r1].cpu(), z[idx - r0, idx + off].cpu()) < 2e-3
    # G = (softmax - onehot) / M; dA = s (G * (1 - w)) B; dD = s G_ii
    G = (z - lse[:, None]).exp() / M
    G[idx - r0, idx + off] -= 1.0 / M
    w1 = 1.0 - ub[off + r0:off + r1] @ ub.T
    ref_da = s * ((G * w1) @ f.float())
    assert rel_err(da[0, r0:r1].cpu(), ref_da.cpu()) < 2e-2
    assert rel_err(ddiag[0, r0:r1].cpu(), (s * G[idx - r0, idx + off]).cpu()) < 2e-2
    # dB for 256 columns: needs G over all M rows, block by block
    c0, c1 = 30000, 30256
    ref_db = torch.zeros(c1 - c0, d, device=dev)
    for i0 in range(0, M, 2048):
        zz = z_rows(i0, i0 + 2048)
        GG = (zz - row_lse[0, i0:i0 + 2048, None]).exp() / M
        ii = torch.arange(i0, i0 + 2048, device=dev)
        GG[ii - i0, ii + off] -= 1.0 / M
        ww = 1.0 - ub[off + i0:off + i0 + 2048] @ ub[c0:c1].T
        ref_db += s * (GG[:, c0:c1] * ww).T @ f[off + i0:off + i0 + 2048].float()
    assert rel_err(db[0, c0:c1].cpu(), ref_db.cpu()) < 2e-2

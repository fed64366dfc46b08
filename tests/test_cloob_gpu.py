"""GPU parity of the CLOOB loss (Hopfield retrieval + InfoLOOB) against the reference's golden
vectors and the CPU oracle.  fp32 path: rel 1e-5 loss, 1e-4 gradients."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import losses_oracle as O

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("case", ["n128_d64", "n200_d96"])
def test_cloob_vs_reference_golden(lib_built, case):
    from multimodal_2025_b200.losses import CLOOBLoss, hopfield
    g = load_golden("cloob.npz")[case]
    x = T(g["x"]).cuda().requires_grad_(True)
    y = T(g["y"]).cuda().requires_grad_(True)
    loss = CLOOBLoss(30.0, 15.0, "cuda")(x, y)
    loss.backward()
    assert rel_err(loss.detach().cpu(), g["loss"]) < 1e-5
    assert rel_err(x.grad.cpu(), g["gx"]) < 1e-4
    assert rel_err(y.grad.cpu(), g["gy"]) < 1e-4
    with torch.no_grad():
        assert rel_err(hopfield(x, x, 15.0).cpu(), g["pxx"]) < 1e-5
        assert rel_err(hopfield(y, x, 15.0).cpu(), g["pxy"]) < 1e-5


@pytest.mark.parametrize("n,d", [(512, 512), (333, 70)])
def test_cloob_vs_oracle(lib_built, n, d):
    from multimodal_2025_b200.losses import cloob_loss
    x, y = O.synth_embeddings(n, d, 8, correlated=True)
    ref, (rx, ry) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0), [x, y])
    xc = x.cuda().requires_grad_(True)
    yc = y.cuda().requires_grad_(True)
    loss = cloob_loss(xc, yc, 30.0, 15.0, "cuda")
    loss.backward()
    assert rel_err(loss.detach().cpu(), ref) < 1e-5
    assert rel_err(xc.grad.cpu(), rx) < 1e-4 and rel_err(yc.grad.cpu(), ry) < 1e-4


def test_infoloob_rows_sum_property(lib_built):
    """size-independent property: the InfoLOOB gradient w.r.t. x has G rows summing to 0
    (softmax over the negatives minus the positive), so sum_i dx_i . 1 relation holds:
    dL/dx = inv_tau * G y with sum_j G_ij = 0  =>  for y = const rows, dx = 0."""
    from multimodal_2025_b200.losses import infoloob_loss
    n, d = 256, 64
    x = torch.randn(n, d, device="cuda", requires_grad=True)
    y = torch.ones(n, d, device="cuda") * 0.1
    infoloob_loss(x, y, None, 30.0).backward()
    assert x.grad.abs().max() < 1e-6


@pytest.mark.parametrize("n,d", [(256, 256), (1024, 512), (768, 256)])
def test_cloob_bf16_tensor_core_vs_oracle(lib_built, n, d):
    """bf16 inputs of an eligible shape: Hopfield retrievals (forward via the P-panel GEMM,
    backward via the dual-accumulator sweep) and InfoLOOB (MASKDIAG on tcgen05) against the fp32
    oracle on the same bf16-rounded inputs; bf16 contract 2e-3 loss / 2e-2 grads."""
    from multimodal_2025_b200 import _lib
    from multimodal_2025_b200.losses import cloob_loss
    x, y = O.synth_embeddings(n, d, 8, correlated=True, dtype=torch.bfloat16)
    ref, (rx, ry) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0),
                                     [x.float(), y.float()])
    xc = x.cuda().requires_grad_(True)
    yc = y.cuda().requires_grad_(True)
    loss = cloob_loss(xc, yc, 30.0, 15.0, "cuda")
    assert _lib.lib().mm25_debug_last_sim_path() == 1   # InfoLOOB forward ran on tcgen05
    loss.backward()
    assert rel_err(loss.detach().float().cpu(), ref) < 2e-3
    assert rel_err(xc.grad.float().cpu(), rx) < 2e-2
    assert rel_err(yc.grad.float().cpu(), ry) < 2e-2


@pytest.mark.parametrize("M,N,d", [(128, 256, 256), (384, 1024, 512)])
def test_hopfield_tc_pieces_vs_dense(lib_built, M, N, d):
    """retrieval forward O = softmax(beta Q K^T) K and backward (dQ, dK) of the tensor-core path
    against dense fp32 autograd on the same bf16 operands (rectangular Q / K)."""
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(M + N)
    nrm = torch.nn.functional.normalize
    q = nrm(torch.randn(M, d, generator=g), dim=-1).cuda().bfloat16()
    k = nrm(torch.randn(N, d, generator=g), dim=-1).cuda().bfloat16()
    beta = 15.0
    qf = q.float().requires_grad_(True)
    kf = k.float().requires_grad_(True)
    o_ref = torch.softmax(beta * qf @ kf.T, dim=-1) @ kf
    d_o = torch.randn(M, d, generator=g).cuda()
    o_ref.backward(d_o)
    sc = torch.tensor([beta], device="cuda")
    lse, _, _, _ = ops.sim_fwd(q.unsqueeze(0), k.unsqueeze(0), None, None, None, sc,
                               ops.SCALE_MUL, 0.0, ops.SIM_PLAIN, 0, False, ops.IMPL_TC)
    o, _, _, _ = ops.sim_bwd(q.unsqueeze(0), k.unsqueeze(0), None, None, None, sc, ops.SCALE_MUL,
                             0.0, ops.SIM_PLAIN, 0, lse, None, 1.0, 0.0, 0.0, 0.0, None, True,
                             False, False, ops.IMPL_TC)
    o = o[0] / beta
    assert rel_err(o.cpu(), o_ref.detach().cpu()) < 1e-2
    delta = ops.rowdot(d_o, o)
    dq, dk = ops.hopfield_bwd_tc(q, k, d_o, sc, lse[0], delta, True, True)
    assert rel_err(dq.cpu(), qf.grad.cpu()) < 2e-2
    assert rel_err(dk.cpu(), kf.grad.cpu()) < 2e-2


def test_cloob_graphed_step_matches_eager(lib_built):
    """graphs.GraphedStep (generic capture of loss + backward) on the bf16 CLOOB path."""
    from multimodal_2025_b200.graphs import GraphedStep
    from multimodal_2025_b200.losses import CLOOBLoss
    crit = CLOOBLoss(30.0, 15.0, "cuda")
    xs = [O.synth_embeddings(512, 256, 30 + i, correlated=True, dtype=torch.bfloat16)
          for i in range(3)]
    x0, y0 = xs[0]
    step = GraphedStep(lambda a, b: crit(a, b), [x0.cuda().requires_grad_(True),
                                                 y0.cuda().requires_grad_(True)])
    for x, y in xs:
        xc, yc = x.cuda().requires_grad_(True), y.cuda().requires_grad_(True)
        ref = crit(xc, yc)
        ref.backward()
        loss, (gx, gy) = step(x.cuda(), y.cuda())
        torch.cuda.synchronize()
        assert torch.equal(loss, ref.detach())
        assert torch.equal(gx, xc.grad) and torch.equal(gy, yc.grad)

"""GPU parity of the AlignCLIP losses (reference AlignCLIP/align_clip/loss.py) through the
C-ABI: golden vectors produced by the reference for the branches it can run on CPU, the oracle
restatement for the semantic branch.  fp32 contract: rel 1e-5 loss / 1e-4 grads."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import losses_oracle as O

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


def run_ours(fn, tensors, dtype=torch.float32):
    leaves = [t.detach().cuda().to(dtype if t.dim() > 0 else torch.float32).requires_grad_(True)
              for t in tensors]
    out = fn(*leaves)
    out.backward()
    return out.detach().float().cpu(), [l.grad.float().cpu() for l in leaves]


def test_nonsemantic_and_openclip_vs_reference_golden(lib_built):
    from multimodal_2025_b200.losses.align_clip_loss import ClipInModalityLoss, ClipLoss
    G = load_golden("alignclip.npz")
    g = G["nosem"]
    x, y, s = T(g["x"]), T(g["y"]), torch.tensor(float(g["s"]))
    crit = ClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=False)
    loss, (gx, gy, gs) = run_ours(lambda a, b, c: crit(a, b, c), [x, y, s])
    assert rel_err(loss, g["loss"]) < 1e-5
    assert rel_err(gx, g["gx"]) < 1e-4 and rel_err(gy, g["gy"]) < 1e-4
    assert rel_err(gs, g["gs"]) < 1e-4
    g2 = G["openclip"]
    loss, (gx, gy, gs) = run_ours(lambda a, b, c: ClipLoss()(a, b, c), [x, y, s])
    assert rel_err(loss, g2["loss"]) < 1e-5
    assert rel_err(gx, g2["gx"]) < 1e-4 and rel_err(gy, g2["gy"]) < 1e-4
    assert rel_err(gs, g2["gs"]) < 1e-4


SEM_CASES = ["sem_text_n64", "sem_both_n64", "sem_image_n96", "sem_text_n160",
             "sem_both_n200_ragged"]


@pytest.mark.parametrize("tag", SEM_CASES)
def test_semantic_branch_vs_reference_golden(lib_built, tag):
    """fp32 CUDA path against the UNMODIFIED reference class run on CPU
    (tests/golden/make_golden_alignsem.py; AlignCLIP/align_clip/loss.py:226-255)."""
    from multimodal_2025_b200.losses.align_clip_loss import ClipInModalityLoss
    g = load_golden("alignclip_sem.npz")[tag]
    x, y, sem = T(g["x"]), T(g["y"]), T(g["sem"]).cuda()
    s = torch.tensor(float(g["s"]))
    st, si = (bool(v) for v in g["flags"])
    alpha, beta = (float(v) for v in g["ab"])
    crit = ClipInModalityLoss(alpha=alpha, beta=beta, nl_semantic_supervision=True,
                              separate_text=st, separate_image=si)
    loss, (gx, gy, gs) = run_ours(lambda a, b, c: crit(a, b, c, semantic_features=sem), [x, y, s])
    assert rel_err(loss, g["loss"]) < 1e-5
    assert rel_err(gx, g["gx"]) < 1e-4 and rel_err(gy, g["gy"]) < 1e-4
    assert rel_err(gs, g["gs"]) < 1e-4
    d = crit(x.cuda(), y.cuda(), s.cuda(), output_dict=True, semantic_features=sem)
    assert rel_err(d["clip_loss"].cpu(), g["clip_loss"]) < 1e-5
    assert rel_err(d["inModality_loss"].cpu(), g["inmod_loss"]) < 1e-5


def test_alignclip_wrapper_vs_reference_golden(lib_built):
    """losses/alignCLIPLoss.py:2-8 itself (always the semantic branch), reference-run golden."""
    from multimodal_2025_b200.losses import AlignCLIPLoss
    G = load_golden("alignclip_sem.npz")
    c, g = G["sem_text_n64"], G["wrapper_n64"]
    sem = T(c["sem"]).cuda()
    w = AlignCLIPLoss()
    loss, (gx, gy, gs) = run_ours(lambda a, b, s_: w.forward(a, b, s_, semantic_features=sem),
                                  [T(c["x"]), T(c["y"]), torch.tensor(float(c["s"]))])
    assert rel_err(loss, g["loss"]) < 1e-5
    assert rel_err(gx, g["gx"]) < 1e-4 and rel_err(gy, g["gy"]) < 1e-4
    assert rel_err(gs, g["gs"]) < 1e-4


@pytest.mark.parametrize("n,d,dsem,sep_img", [(64, 32, 16, False), (200, 96, 48, False),
                                              (257, 64, 768, True), (512, 512, 768, False)])
def test_semantic_branch_vs_oracle(lib_built, n, d, dsem, sep_img):
    from multimodal_2025_b200.losses.align_clip_loss import ClipInModalityLoss
    x, y = O.synth_embeddings(n, d, 3, correlated=True)
    s = torch.tensor(14.285)
    sem = torch.randn(n, dsem, generator=torch.Generator().manual_seed(1))
    ref, (rx, ry, rs) = O.loss_and_grads(
        lambda a, b, c: O.inmodality_loss(a, b, c, sem, nl_semantic_supervision=True,
                                          separate_image=sep_img)[0], [x, y, s])
    crit = ClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=True,
                              separate_image=sep_img)
    semc = sem.cuda()
    loss, (gx, gy, gs) = run_ours(lambda a, b, c: crit(a, b, c, semantic_features=semc),
                                  [x, y, s])
    assert rel_err(loss, ref) < 1e-5
    assert rel_err(gx, rx) < 1e-4 and rel_err(gy, ry) < 1e-4
    assert rel_err(gs, rs) < 1e-4


def test_alignclip_wrapper_and_dict(lib_built):
    from multimodal_2025_b200.losses import AlignCLIPLoss
    x, y = O.synth_embeddings(96, 64, 5, correlated=True)
    sem = torch.randn(96, 24, generator=torch.Generator().manual_seed(2))
    s = torch.tensor(20.0)
    tot, clip, inm = O.inmodality_loss(x, y, s, sem, nl_semantic_supervision=True)
    w = AlignCLIPLoss()
    assert not callable(w)  # quirk D.2: plain class, only .forward
    out = w.forward(x.cuda(), y.cuda(), s.cuda(), semantic_features=sem.cuda())
    assert rel_err(out.cpu(), tot) < 1e-5
    d = w.criterion(x.cuda(), y.cuda(), s.cuda(), output_dict=True, semantic_features=sem.cuda())
    assert set(d) == {"total_loss", "clip_loss", "inModality_loss"}
    assert rel_err(d["clip_loss"].cpu(), clip) < 1e-5
    assert rel_err(d["inModality_loss"].cpu(), inm) < 1e-5


def test_alignclip_bf16_tensor_core_cross_modal(lib_built):
    """bf16 inputs: the cross-modal term runs the tcgen05 kernels (2e-3 / 2e-2 contract)."""
    from multimodal_2025_b200.losses.align_clip_loss import ClipLoss
    x, y = O.synth_embeddings(1024, 256, 7, correlated=True, dtype=torch.bfloat16)
    s = torch.tensor(30.0)
    ref, (rx, ry, rs) = O.loss_and_grads(O.openclip_loss, [x.float(), y.float(), s])
    loss, (gx, gy, gs) = run_ours(lambda a, b, c: ClipLoss()(a, b, c), [x, y, s], torch.bfloat16)
    assert rel_err(loss, ref) < 2e-3
    assert rel_err(gx, rx) < 2e-2 and rel_err(gy, ry) < 2e-2
    assert rel_err(gs, rs) < 2e-2


@pytest.mark.parametrize("n,d,dsem,sep_img", [(256, 256, 128, False), (512, 512, 768, True),
                                              (1024, 256, 64, False)])
def test_semantic_branch_bf16_tensor_core(lib_built, n, d, dsem, sep_img):
    """bf16 features of an eligible shape: the in-modality semantic term runs the two-accumulator
    tcgen05 kernels (csrc/sim_tc_mod.cu); oracle fed the same bf16-rounded features, bf16
    contract (2e-3 loss / 2e-2 grads)."""
    from multimodal_2025_b200.losses.align_clip_loss import ClipInModalityLoss
    x, y = O.synth_embeddings(n, d, 3, correlated=True, dtype=torch.bfloat16)
    s = torch.tensor(14.285)
    sem = torch.randn(n, dsem, generator=torch.Generator().manual_seed(1))
    ref, (rx, ry, rs) = O.loss_and_grads(
        lambda a, b, c: O.inmodality_loss(a, b, c, sem, nl_semantic_supervision=True,
                                          separate_image=sep_img)[0],
        [x.float(), y.float(), s])
    crit = ClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=True,
                              separate_image=sep_img)
    semc = sem.cuda()
    loss, (gx, gy, gs) = run_ours(lambda a, b, c: crit(a, b, c, semantic_features=semc),
                                  [x, y, s], torch.bfloat16)
    assert rel_err(loss, ref) < 2e-3
    assert rel_err(gx, rx) < 2e-2 and rel_err(gy, ry) < 2e-2
    assert rel_err(gs, rs) < 2e-2


@pytest.mark.parametrize("M,N,d,d2,off", [(128, 256, 256, 64, 0), (256, 1024, 512, 192, 512),
                                          (384, 768, 256, 768, 384)])
def test_modulated_tc_matches_simt(lib_built, M, N, d, d2, off):
    """C-ABI level: MODULATED forward statistics and backward (da, db, dscale, ddiag) of the
    tcgen05 path against the fp32 SIMT kernels on the same bf16 operands, rectangular blocks
    with shifted positives (the row-sharded layout)."""
    from multimodal_2025_b200 import ops
    g = torch.Generator().manual_seed(M + N)
    nrm = torch.nn.functional.normalize
    a = nrm(torch.randn(1, M, d, generator=g), dim=-1).cuda().bfloat16()
    b = nrm(torch.randn(1, N, d, generator=g), dim=-1).cuda().bfloat16()
    v = nrm(torch.randn(1, N, d2, generator=g), dim=-1).cuda()
    u = v[:, off:off + M].contiguous()
    # semantic features rounded to bf16 once so that both paths see the same operands
    u, v = u.bfloat16().float(), v.bfloat16().float()
    D = torch.rand(1, M, generator=g).cuda()
    sc = torch.tensor([14.285], device="cuda")
    outs = {}
    for impl in (ops.IMPL_SIMT, ops.IMPL_AUTO):
        row_lse, _, _, diag = ops.sim_fwd(a, b, u, v, D, sc, ops.SCALE_MUL, 0.0,
                                          ops.SIM_MODULATED, off, False, impl)
        gout = torch.tensor([0.7], device="cuda")
        bw = ops.sim_bwd(a, b, u, v, D, sc, ops.SCALE_MUL, 0.0, ops.SIM_MODULATED, off, row_lse,
                         None, 1.0 / M, 0.0, 1.0 / M, 0.0, gout, True, True, True, impl)
        outs[impl] = (row_lse, diag) + tuple(bw)
        from multimodal_2025_b200 import _lib
        assert _lib.lib().mm25_debug_last_sim_path() == (0 if impl == ops.IMPL_SIMT else 2)
    ref, got = outs[ops.IMPL_SIMT], outs[ops.IMPL_AUTO]
    assert (got[0] - ref[0]).abs().max().item() < 2e-4          # row LSE (fp32 accumulate)
    assert (got[1] - ref[1]).abs().max().item() < 2e-4 * 15      # positives
    for k in (2, 3):                                             # da, db (bf16 G'' panel)
        assert (got[k] - ref[k]).norm().item() <= 2e-2 * ref[k].norm().item(), k
    assert abs(got[4].item() - ref[4].item()) <= 2e-2 * abs(ref[4].item()) + 1e-6
    assert (got[5] - ref[5]).norm().item() <= 1e-3 * ref[5].norm().item() + 1e-7


@pytest.mark.parametrize("dtype,k,b,d,tl,tg", [(torch.float32, 4, 96, 64, 1e-5, 1e-4),
                                               (torch.bfloat16, 4, 256, 256, 2e-3, 2e-2),
                                               (torch.bfloat16, 8, 256, 512, 2e-3, 2e-2)])
@pytest.mark.parametrize("need_scale", [True, False])
def test_cached_negatives_live_rows_vs_oracle(lib_built, dtype, k, b, d, tl, tg, need_scale):
    """SURVEY §8f rank 4 (AlignCLIP/main/train.py:139-157): micro-step j of the accumulation loop;
    loss of the spliced batch (oracle.openclip_loss) and gradients of the live rows only, fp32
    SIMT and bf16 tcgen05 rectangular kernels (positives at column offset j b)."""
    from multimodal_2025_b200.losses.negatives_cache import cached_negatives_clip_loss
    j = k - 2
    x, y = O.synth_embeddings(k * b, d, 12, correlated=True, dtype=dtype)
    g = torch.Generator().manual_seed(3)
    live = slice(j * b, (j + 1) * b)
    li = torch.nn.functional.normalize(x[live].float() + 0.05 * torch.randn(b, d, generator=g),
                                       dim=-1).to(dtype)
    lt = torch.nn.functional.normalize(y[live].float() + 0.05 * torch.randn(b, d, generator=g),
                                       dim=-1).to(dtype)
    a = li.float().clone().requires_grad_(True)
    t = lt.float().clone().requires_grad_(True)
    s = torch.tensor(14.285, requires_grad=need_scale)
    xi = torch.cat([x[:j * b].float(), a, x[(j + 1) * b:].float()])
    xt = torch.cat([y[:j * b].float(), t, y[(j + 1) * b:].float()])
    ref = O.openclip_loss(xi, xt, s)
    ref.backward()
    ci = [c.cuda() for c in x.split(b)]
    ct = [c.cuda() for c in y.split(b)]
    a2 = li.cuda().requires_grad_(True)
    t2 = lt.cuda().requires_grad_(True)
    s2 = torch.tensor(14.285, device="cuda", requires_grad=need_scale)
    loss = cached_negatives_clip_loss(a2, t2, s2, ci, ct, j)
    loss.backward()
    assert rel_err(loss.detach().cpu(), ref.detach()) < tl
    assert rel_err(a2.grad.float().cpu(), a.grad) < tg
    assert rel_err(t2.grad.float().cpu(), t.grad) < tg
    if need_scale:
        assert rel_err(s2.grad.cpu(), s.grad) < tg

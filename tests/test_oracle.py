"""CPU: the oracle restatement against golden vectors produced by the reference itself
(tests/golden/make_golden.py).  Tolerances are fp32 re-association noise only."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import losses_oracle as O


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("case", ["n256_d512", "n256_d512_clamp", "n192_d96_eps0", "n70_d40_raw",
                                  "n96_d64_nonorm"])
def test_clip_forward_matches_reference(case):
    g = load_golden("clip_forward.npz")[case]
    norm = case != "n96_d64_nonorm"
    loss, (gx, gy, gl) = O.loss_and_grads(
        lambda a, b, c: O.clip_loss(a, b, c, float(g["eps"]), norm),
        [T(g["x"]), T(g["y"]), torch.tensor(float(g["ls"]))])
    assert rel_err(loss, g["loss"]) < 1e-6
    assert rel_err(gx, g["gx"]) < 1e-5 and rel_err(gy, g["gy"]) < 1e-5
    assert abs(float(gl) - float(g["gl"])) <= 1e-5 * max(1.0, abs(float(g["gl"])))


def test_clip_fp32_error_budget_vs_fp64():
    """How far the fp32 reference itself is from fp64 truth: sets what 'rel 1e-5' can mean."""
    x, y = O.synth_embeddings(256, 512, 0)
    ls = torch.tensor(2.6592)
    l32 = O.clip_loss(x, y, ls)
    l64 = O.clip_loss(x.double(), y.double(), ls.double(), dtype=torch.float64)
    assert rel_err(l32, l64) < 2e-6


@pytest.mark.parametrize("key", ["eps0.1_ls2.6592", "eps0.1_ls4.2", "eps0.0_ls2.6592",
                                 "eps0.0_ls4.2"])
def test_clip_sampled_matches_reference(key):
    G = load_golden("clip_sampled.npz")
    g, top = G[key], G[""]
    eps = float(key.split("_")[0][3:])
    ls = float(key.split("ls")[1])
    w = T(top["w"])
    loss, (gx, gy, gl) = O.loss_and_grads(
        lambda a, b, c: O.clip_loss_sampled(a, b, c, eps) * w,
        [T(top["x"]), T(top["y"]), torch.tensor(ls)])
    assert rel_err(loss / w, g["loss"]) < 1e-6
    assert rel_err(gx, g["gx"]) < 1e-5 and rel_err(gy, g["gy"]) < 1e-5
    assert abs(float(gl) - float(g["gl"])) <= 2e-5 * max(1.0, abs(float(g["gl"])))


@pytest.mark.parametrize("case", ["n128_d64", "n200_d96"])
def test_cloob_matches_reference(case):
    g = load_golden("cloob.npz")[case]
    x, y = T(g["x"]), T(g["y"])
    loss, (gx, gy) = O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0), [x, y])
    assert rel_err(loss, g["loss"]) < 1e-6
    assert rel_err(gx, g["gx"]) < 1e-4 and rel_err(gy, g["gy"]) < 1e-4
    assert rel_err(O.hopfield(x, x, 15.0), g["pxx"]) < 1e-6
    assert rel_err(O.hopfield(y, x, 15.0), g["pxy"]) < 1e-6


def test_alignclip_nonsemantic_and_openclip_match_reference():
    G = load_golden("alignclip.npz")
    g = G["nosem"]
    x, y, s = T(g["x"]), T(g["y"]), torch.tensor(float(g["s"]))
    loss, (gx, gy, gs) = O.loss_and_grads(
        lambda a, b, c: O.inmodality_loss(a, b, c)[0], [x, y, s])
    assert rel_err(loss, g["loss"]) < 1e-6
    assert rel_err(gx, g["gx"]) < 1e-5 and rel_err(gy, g["gy"]) < 1e-5
    assert rel_err(gs, g["gs"]) < 1e-4
    g2 = G["openclip"]
    loss, (gx, gy, gs) = O.loss_and_grads(O.openclip_loss, [x, y, s])
    assert rel_err(loss, g2["loss"]) < 1e-6 and rel_err(gx, g2["gx"]) < 1e-5


def test_alignclip_semantic_branch_consistency():
    """The semantic branch cannot run on CPU in the reference (loss.py:233,237). Its pieces are
    pinned indirectly: with semantic_sim == 1 off-diagonal it must reduce to known terms."""
    x, y = O.synth_embeddings(64, 32, 3, correlated=True)
    s = torch.tensor(10.0)
    sem = torch.randn(64, 16, generator=torch.Generator().manual_seed(0))
    tot, clip, inm = O.inmodality_loss(x, y, s, sem, nl_semantic_supervision=True)
    # manual: Z_T = s * [(T T^T) o (1 - S S^T) + diag(i.t)]
    sh = sem / sem.norm(dim=-1, keepdim=True)
    zt = s * ((y @ y.T) * (1 - sh @ sh.T) + torch.diag((x * y).sum(-1)))
    ref = 0.5 * torch.nn.functional.cross_entropy(zt, torch.arange(64))
    assert rel_err(inm, ref) < 1e-6
    assert rel_err(clip, 2 * O.openclip_loss(x, y, s)) < 1e-6


SEM_CASES = ["sem_text_n64", "sem_both_n64", "sem_image_n96", "sem_text_n160",
             "sem_both_n200_ragged"]


@pytest.mark.parametrize("tag", SEM_CASES)
def test_alignclip_semantic_oracle_matches_reference_golden(tag):
    """the SEMANTIC branch (loss.py:226-255) pinned by the UNMODIFIED reference class run on CPU
    (tests/golden/make_golden_alignsem.py patches Tensor.get_device for the call only)."""
    g = load_golden("alignclip_sem.npz")[tag]
    x, y, sem = (torch.from_numpy(g[k]) for k in ("x", "y", "sem"))
    s = torch.tensor(float(g["s"]))
    st, si = (bool(v) for v in g["flags"])
    alpha, beta = (float(v) for v in g["ab"])
    kw = dict(alpha=alpha, beta=beta, nl_semantic_supervision=True, separate_text=st,
              separate_image=si)
    loss, (gx, gy, gs) = O.loss_and_grads(
        lambda a, b, c: O.inmodality_loss(a, b, c, sem, **kw)[0], [x, y, s])
    assert rel_err(loss, g["loss"]) < 1e-6
    assert rel_err(gx, g["gx"]) < 1e-5 and rel_err(gy, g["gy"]) < 1e-5
    assert rel_err(gs, g["gs"]) < 1e-5
    tot, clip, inm = O.inmodality_loss(x, y, s, sem, **kw)
    assert rel_err(clip, g["clip_loss"]) < 1e-6 and rel_err(inm, g["inmod_loss"]) < 1e-6


@pytest.mark.parametrize("n,d,eps,nslab", [(256, 64, 0.1, 4), (96, 32, 0.0, 3), (130, 48, 0.1, 5)])
def test_clip_loss_slabs_combine_to_the_pinned_clip_loss(n, d, eps, nslab):
    """bench.py times `clip_loss_slab` (a bounded sample of a batch too big for the host); its
    pieces over a partition of the rows must give exactly the golden-pinned `clip_loss`, value
    and gradients."""
    x, y = O.synth_embeddings(n, d, 4, correlated=True)
    x = x * 1.7
    ls = torch.tensor(2.9)
    ref, (rx, ry, rl) = O.loss_and_grads(O.clip_loss, [x, y, ls], eps, True)
    bounds = [round(i * n / nslab) for i in range(nslab + 1)]

    def slabbed(a, b, c):
        parts = [O.clip_loss_slab(a, b, c, bounds[i], bounds[i + 1] - bounds[i], eps, True)
                 for i in range(nslab)]
        return O.clip_loss_from_slabs(parts, n, eps)

    got, (gx, gy, gl) = O.loss_and_grads(slabbed, [x, y, ls])
    assert rel_err(got, ref) < 1e-6
    assert rel_err(gx, rx) < 1e-5 and rel_err(gy, ry) < 1e-5 and rel_err(gl, rl) < 1e-5

"""GPU parity of the vCLIP path through the C-ABI: sampler kernels in noise-injection mode
against the reference's golden vectors (tests/golden/dist.npz) and the CPU oracle, the device
log-Bessel against mpmath, the Philox path statistically.  fp32 contract: rel 1e-5 values /
1e-4 gradients."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import dist_oracle as D

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


def C(a):
    return T(a).cuda()


@pytest.mark.parametrize("case", ["vmf_b6_d16", "vmf_b3_d512"])
def test_vmf_samples_and_loc_grad_vs_reference(lib_built, case):
    from multimodal_2025_b200.losses.distributions import VonMisesFisher
    g = load_golden("dist.npz")[case]
    loc = C(g["loc"]).requires_grad_(True)
    dist = VonMisesFisher(loc, C(g["kappa"]))
    N = g["samples"].shape[0]
    dist._injected_noise = {"v": C(g["v"]).permute(1, 0, 2).contiguous(), "eps": C(g["eps"]),
                            "u": C(g["u"])}
    s = dist.rsample((N,))
    assert s.shape == g["samples"].shape
    assert rel_err(s.detach().cpu(), g["samples"]) < 1e-5
    (s * C(g["W"])).sum().backward()
    assert rel_err(loc.grad.cpu(), g["gloc"]) < 1e-4


@pytest.mark.parametrize("case", ["vmf_b6_d16", "vmf_b3_d512"])
def test_vmf_entropy_logprob_vs_reference(lib_built, case):
    from multimodal_2025_b200.losses.distributions import VonMisesFisher
    g = load_golden("dist.npz")[case]
    kappa = C(g["kappa"]).requires_grad_(True)
    dist = VonMisesFisher(C(g["loc"]), kappa)
    ent = dist.entropy()
    ent.sum().backward()
    assert rel_err(ent.detach().cpu(), g["entropy"]) < 1e-5
    assert rel_err(kappa.grad.cpu(), g["gkappa"]) < 1e-4
    lp = VonMisesFisher(C(g["loc"]), C(g["kappa"])).log_prob(C(g["samples"])[0])
    assert rel_err(lp.cpu(), g["log_prob"]) < 1e-5


@pytest.mark.parametrize("case,d", [("logc_d512", 512), ("logc_d3", 3)])
def test_log_partition_vs_reference_mpmath(lib_built, case, d):
    from multimodal_2025_b200.losses.distributions import vMFLogPartition
    g = load_golden("dist.npz")[case]
    kappa = C(g["kappa"]).requires_grad_(True)
    logC = vMFLogPartition.apply(d, kappa)
    logC.sum().backward()
    assert rel_err(logC.detach().cpu(), g["logC"]) < 1e-6
    # the reference's derivative is exp of a difference of fp32-rounded logs (VonMisesFisher.py:60)
    assert torch.allclose(kappa.grad.cpu(), T(g["dlogC"]), rtol=2e-2, atol=1e-6)


@pytest.mark.parametrize("nu", [0.0, 0.5, 3.0, 15.5, 16.0, 63.0, 255.0, 511.0])
def test_log_bessel_vs_mpmath(lib_built, nu):
    from multimodal_2025_b200 import ops
    kappa = torch.tensor([1e-3, 0.3, 1.0, 7.0, 30.0, 150.0, 199.0, 201.0, 1e3, 1e4, 1e6],
                         dtype=torch.float64)
    ref0 = D.log_bessel_mpmath(nu, kappa)
    ref1 = D.log_bessel_mpmath(nu + 1, kappa)
    o0, o1 = ops.log_bessel(kappa.cuda(), nu)
    err0 = ((o0.cpu() - ref0).abs() / ref0.abs().clamp_min(1.0)).max()
    err1 = ((o1.cpu() - ref1).abs() / ref1.abs().clamp_min(1.0)).max()
    assert err0 < 1e-8 and err1 < 1e-8, (float(err0), float(err1))


def test_projected_normal_vs_reference(lib_built):
    from multimodal_2025_b200.losses.distributions import ProjectedNormal
    g = load_golden("dist.npz")["pn"]
    mu, sg = C(g["mu"]).requires_grad_(True), C(g["sigma"]).requires_grad_(True)
    pn = ProjectedNormal(mu, sg)
    pn._injected_noise = {"eps": C(g["eps"])}
    s = pn.rsample((g["samples"].shape[0],))
    assert rel_err(s.detach().cpu(), g["samples"]) < 1e-5
    (s * C(g["W"])).sum().backward()
    assert rel_err(mu.grad.cpu(), g["gmu"]) < 1e-4 and rel_err(sg.grad.cpu(), g["gsigma"]) < 1e-4
    assert rel_err(pn.entropy().detach().cpu(), g["entropy"]) < 1e-6
    assert rel_err(pn.loc.detach().cpu(), g["loc"]) < 1e-6


@pytest.mark.parametrize("B,d,S", [(5, 16, 6), (4, 512, 3), (7, 3, 5), (3, 130, 4)])
def test_power_spherical_vs_oracle(lib_built, B, d, S):
    """unpinned restatement: same draws -> same samples and gradients (incl. the kappa gradient
    through the Beta draw: torch's implicit reparameterisation vs mm25_beta_grad)."""
    from multimodal_2025_b200.losses.distributions import PowerSpherical
    gen = torch.Generator().manual_seed(B * d)
    loc = torch.nn.functional.normalize(torch.randn(B, d, generator=gen), dim=-1)
    scale = torch.exp(torch.rand(B, generator=gen) * 4 + 1)
    W = torch.randn(S, B, d, generator=gen)
    v = torch.randn(S, B, d - 1, generator=gen)
    lo, sc = loc.clone().requires_grad_(True), scale.clone().requires_grad_(True)
    a, b = D.ps_alpha_beta(d, sc)
    torch.manual_seed(1)
    z = torch.distributions.Beta(a, b).rsample((S,))
    y = D.ps_rsample(lo, sc, z, v)
    (y * W).sum().backward()
    lc, scc = loc.cuda().requires_grad_(True), scale.cuda().requires_grad_(True)
    ps = PowerSpherical(lc, scc)
    ps._injected_noise = {"z": z.detach().cuda(), "v": v.cuda()}
    yc = ps.rsample((S,))
    assert rel_err(yc.detach().cpu(), y.detach()) < 1e-5
    (yc * W.cuda()).sum().backward()
    assert rel_err(lc.grad.cpu(), lo.grad) < 1e-4
    assert rel_err(scc.grad.cpu(), sc.grad) < 2e-4
    # torch lgamma / digamma differences in fp32 cancel poorly at small d: absolute tolerance
    assert torch.allclose(ps.entropy().detach().cpu(), D.ps_entropy(d, scale), rtol=1e-4, atol=1e-4)


def test_beta_grad_all_branches_vs_torch(lib_built):
    from multimodal_2025_b200 import ops
    cases = [(0.01, 0.8, 1.2), (0.3, 2.0, 3.0), (0.97, 1.5, 0.7), (0.5, 30.0, 40.0),
             (0.428, 30.0, 40.0), (0.9, 300.0, 255.5), (0.6, 265.5, 255.5), (0.999, 8.0, 9.0),
             (0.2, 0.5, 0.5)]
    z = torch.tensor([[c[0] for c in cases]])
    a = torch.tensor([c[1] for c in cases])
    b = torch.tensor([c[2] for c in cases])
    x = torch.stack([z[0], 1 - z[0]], -1)
    conc = torch.stack([a, b], -1)
    ref = torch._dirichlet_grad(x, conc, conc.sum(-1, True).expand_as(conc))[:, 0] * (1 - z[0])
    got = ops.beta_grad(z.cuda(), a.cuda(), b.cuda()).cpu()[0]
    assert torch.allclose(got, ref, rtol=2e-4, atol=1e-7), (got, ref)


@pytest.mark.parametrize("eps_ls", [0.1, 0.0])
def test_vclip_vmf_end_to_end_vs_reference(lib_built, eps_ls):
    from multimodal_2025_b200.losses.distributions import VonMisesFisher
    from multimodal_2025_b200.losses.vclipLoss import VClipLoss
    g = load_golden("dist.npz")["vclip_vmf_eps%g" % eps_ls]
    leaves = [C(g[k]).clone().requires_grad_(True) for k in ("li", "lt", "ki", "kt", "ls")]
    di, dt_ = VonMisesFisher(leaves[0], leaves[2]), VonMisesFisher(leaves[1], leaves[3])
    di._injected_noise = {"v": C(g["vi"]).permute(1, 0, 2).contiguous(), "eps": C(g["ei"]),
                          "u": C(g["ui"])}
    dt_._injected_noise = {"v": C(g["vt"]).permute(1, 0, 2).contiguous(), "eps": C(g["et"]),
                           "u": C(g["ut"])}
    S = g["vi"].shape[1]
    res = VClipLoss(num_samples=S, label_smoothing=eps_ls).cuda()(
        di, dt_, 1.0 / leaves[2], 1.0 / leaves[3], leaves[4], True)
    assert set(res) == {"total_loss", "clip_loss", "image_kl_loss", "var_reg", "text_kl_loss"}
    res["total_loss"].backward()
    for k in res:
        assert rel_err(res[k].detach().cpu(), g[k]) < 1e-5, k
    for n, l in zip(("li", "lt", "ki", "kt", "ls"), leaves):
        assert rel_err(l.grad.cpu(), g["g" + n]) < 1e-4, n


def test_vclip_projected_normal_and_eval_vs_reference(lib_built):
    from multimodal_2025_b200.losses.distributions import ProjectedNormal, VonMisesFisher
    from multimodal_2025_b200.losses.vclipLoss import VClipLoss
    G = load_golden("dist.npz")
    g = G["vclip_pn"]
    leaves = [C(g[k]).clone().requires_grad_(True) for k in ("mi", "mt", "si", "st", "ls")]
    di, dt_ = ProjectedNormal(leaves[0], leaves[2]), ProjectedNormal(leaves[1], leaves[3])
    di._injected_noise, dt_._injected_noise = {"eps": C(g["ei"])}, {"eps": C(g["et"])}
    res = VClipLoss(num_samples=g["ei"].shape[0], label_smoothing=0.1).cuda()(
        di, dt_, leaves[2].mean(-1), leaves[3].mean(-1), leaves[4], True)
    res["total_loss"].backward()
    for k in res:
        assert rel_err(res[k].detach().cpu(), g[k]) < 1e-5, k
    for n, l in zip(("mi", "mt", "si", "st", "ls"), leaves):
        assert rel_err(l.grad.cpu(), g["g" + n]) < 1e-4, n
    e = G["vclip_vmf_eval"]
    res = VClipLoss(label_smoothing=0.1).cuda()(
        VonMisesFisher(C(e["li"]), C(e["ki"])), VonMisesFisher(C(e["lt"]), C(e["kt"])),
        1.0 / C(e["ki"]), 1.0 / C(e["kt"]), torch.tensor(2.6592).cuda(), False)
    for k in res:
        assert rel_err(res[k].detach().cpu(), e[k]) < 1e-5, k


def test_philox_samplers_statistics(lib_built):
    """in-kernel RNG: unit norm, E[mu.x] of PowerSpherical = (a-b)/(a+b), vMF mean resultant
    length close to A_d(kappa), ProjectedNormal mean direction, determinism under manual_seed."""
    from multimodal_2025_b200.losses.distributions import (PowerSpherical, ProjectedNormal,
                                                           VonMisesFisher)
    B, d, S = 16, 64, 2000
    gen = torch.Generator().manual_seed(3)
    loc = torch.nn.functional.normalize(torch.randn(B, d, generator=gen), dim=-1).cuda()
    kappa = torch.linspace(5, 400, B).cuda()
    torch.manual_seed(7)
    y = PowerSpherical(loc, kappa).rsample((S,))
    assert (y.norm(dim=-1) - 1).abs().max() < 1e-4
    a, b = (d - 1) / 2 + kappa, (d - 1) / 2
    assert torch.allclose((y * loc).sum(-1).mean(0), (a - b) / (a + b), atol=0.01)
    torch.manual_seed(7)
    y2 = PowerSpherical(loc, kappa).rsample((S,))
    assert torch.equal(y, y2)
    x = VonMisesFisher(loc, kappa).rsample((S,))
    assert (x.norm(dim=-1) - 1).abs().max() < 1e-4
    from oracle.dist_oracle import log_bessel_mpmath
    nu = d / 2 - 1
    A = (log_bessel_mpmath(nu + 1, kappa.cpu()) - log_bessel_mpmath(nu, kappa.cpu())).exp()
    assert torch.allclose((x * loc).sum(-1).mean(0).cpu().double(), A, atol=0.01)
    p = ProjectedNormal(loc * 3.0, torch.full((B, d), 0.05).cuda()).rsample((S,))
    assert (p.norm(dim=-1) - 1).abs().max() < 1e-4
    assert ((p.mean(0) * loc).sum(-1) > 0.98).all()

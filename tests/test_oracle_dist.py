"""CPU: oracle/dist_oracle.py against golden vectors produced by the reference itself
(tests/golden/make_golden_dist.py), and self-checks of the unpinned PowerSpherical restatement."""
import math

import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import dist_oracle as D


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("case", ["vmf_b6_d16", "vmf_b3_d512"])
def test_vmf_sample_replay_matches_reference(case):
    g = load_golden("dist.npz")[case]
    loc = T(g["loc"]).requires_grad_(True)
    s = D.vmf_sample(D.vmf_normalize_loc(loc), T(g["kappa"]), T(g["v"]), T(g["eps"]), T(g["u"]))
    assert rel_err(s, g["samples"]) < 1e-6
    (s * T(g["W"])).sum().backward()
    assert rel_err(loc.grad, g["gloc"]) < 1e-5


@pytest.mark.parametrize("case", ["vmf_b6_d16", "vmf_b3_d512"])
def test_vmf_entropy_and_logprob_match_reference(case):
    g = load_golden("dist.npz")[case]
    d = g["loc"].shape[-1]
    kappa = T(g["kappa"]).requires_grad_(True)
    ent = D.vmf_entropy(d, kappa)
    ent.sum().backward()
    assert rel_err(ent, g["entropy"]) < 1e-6
    assert rel_err(kappa.grad, g["gkappa"]) < 1e-5
    lp = D.vmf_log_prob(D.vmf_normalize_loc(T(g["loc"])), T(g["kappa"]), T(g["samples"])[0])
    assert rel_err(lp, g["log_prob"]) < 1e-6


def test_projected_normal_matches_reference():
    g = load_golden("dist.npz")["pn"]
    mu, sg = T(g["mu"]).requires_grad_(True), T(g["sigma"]).requires_grad_(True)
    s = D.pn_rsample(mu, sg, T(g["eps"]))
    assert rel_err(s, g["samples"]) < 1e-6
    (s * T(g["W"])).sum().backward()
    assert rel_err(mu.grad, g["gmu"]) < 1e-5 and rel_err(sg.grad, g["gsigma"]) < 1e-5
    assert rel_err(D.pn_entropy(T(g["mu"]), T(g["sigma"])), g["entropy"]) < 1e-6


@pytest.mark.parametrize("eps_ls", [0.1, 0.0])
def test_vclip_vmf_matches_reference(eps_ls):
    g = load_golden("dist.npz")["vclip_vmf_eps%g" % eps_ls]
    li, lt = T(g["li"]).requires_grad_(True), T(g["lt"]).requires_grad_(True)
    ki, kt = T(g["ki"]).requires_grad_(True), T(g["kt"]).requires_grad_(True)
    ls = T(g["ls"]).clone().requires_grad_(True)
    d = li.shape[-1]
    si = D.vmf_sample(D.vmf_normalize_loc(li), ki, T(g["vi"]), T(g["ei"]), T(g["ui"]))
    st = D.vmf_sample(D.vmf_normalize_loc(lt), kt, T(g["vt"]), T(g["et"]), T(g["ut"]))
    res = D.vclip_total(si, st, D.vmf_kl(d, ki), D.vmf_kl(d, kt), 1.0 / ki, 1.0 / kt, ls,
                        label_smoothing=eps_ls)
    res["total_loss"].backward()
    for k in ("total_loss", "clip_loss", "image_kl_loss", "var_reg", "text_kl_loss"):
        assert rel_err(res[k], g[k]) < 1e-5, k
    assert rel_err(li.grad, g["gli"]) < 1e-4 and rel_err(lt.grad, g["glt"]) < 1e-4
    assert rel_err(ki.grad, g["gki"]) < 1e-4 and rel_err(kt.grad, g["gkt"]) < 1e-4
    assert rel_err(ls.grad, g["gls"]) < 1e-4


def test_power_spherical_self_checks():
    """unpinned restatement: unit norm, E[loc.x] = (a-b)/(a+b), KL -> 0 as kappa -> 0."""
    torch.manual_seed(0)
    B, d, S = 4, 32, 4000
    loc = torch.nn.functional.normalize(torch.randn(B, d), dim=-1)
    scale = torch.tensor([1.0, 10.0, 50.0, 200.0])
    a, b = D.ps_alpha_beta(d, scale)
    z = torch.distributions.Beta(a, b).sample((S,))
    y = D.ps_rsample(loc, scale, z, torch.randn(S, B, d - 1))
    assert (y.norm(dim=-1) - 1).abs().max() < 1e-4
    m = (y * loc).sum(-1).mean(0)
    assert torch.allclose(m, (a - b) / (a + b), atol=0.02)
    assert abs(float(D.ps_kl_uniform(d, torch.tensor([1e-4])))) < 1e-3
    assert float(D.ps_kl_uniform(d, torch.tensor([50.0]))) > 0
    # entropy of the uniform = log surface area of S^{d-1}
    assert abs(D.hu_entropy(3) - math.log(4 * math.pi)) < 1e-9

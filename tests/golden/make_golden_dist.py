"""Golden vectors of the vCLIP path, produced by RUNNING THE REFERENCE ITSELF
(losses/distributions/VonMisesFisher.py, ProjectedNormal.py, losses/vclipLoss.py imported from
/root/reference; build container only).  Writes tests/golden/dist.npz.

The reference draws its noise from torch's global generator.  To make the draws available to
the CUDA path (noise-injection mode) each case is run under torch.manual_seed(k) and the exact
same sequence of RNG calls is then replayed under the same seed to capture the noise.
`power_spherical` is absent here; a shim built on oracle/dist_oracle.py is injected so that
vclipLoss.py imports (its PowerSpherical branch is NOT used for goldens: parity unpinned).
"""
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("MM25_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(OUT))
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

shim = types.ModuleType("power_spherical")
shim.PowerSpherical = type("PowerSpherical", (), {})
shim.HypersphericalUniform = type("HypersphericalUniform", (), {})
sys.modules["power_spherical"] = shim

from losses.distributions.VonMisesFisher import VonMisesFisher, vMFLogPartition  # noqa: E402
from losses.distributions.ProjectedNormal import ProjectedNormal  # noqa: E402
from losses.vclipLoss import VClipLoss  # noqa: E402


def replay_vmf(seed, B, N, d):
    """the RNG calls of VonMisesFisher._sample_impl (:180-222), first rejection round."""
    torch.manual_seed(seed)
    vs, es, us = [], [], []
    for _ in range(B):
        vs.append(torch.randn(N, d - 1))
        beta = torch.distributions.Beta(torch.tensor(0.5 * (d - 1)), torch.tensor(0.5 * (d - 1)))
        uni = torch.distributions.Uniform(torch.tensor(0.0), torch.tensor(1.0))
        es.append(beta.sample([10 * N]))
        us.append(uni.sample([10 * N]))
    return torch.stack(vs), torch.stack(es), torch.stack(us)


def replay_pn(seed, shape):
    torch.manual_seed(seed)
    return torch.distributions.utils._standard_normal(shape, dtype=torch.float32, device="cpu")


def main():
    torch.set_num_threads(4)
    out = {}
    g = torch.Generator().manual_seed(5)

    # ---- vMF: samples + grad to loc, log_prob, entropy + grad to kappa -----------------------
    for tag, (B, d, N) in {"vmf_b6_d16": (6, 16, 5), "vmf_b3_d512": (3, 512, 4)}.items():
        loc = torch.randn(B, d, generator=g) * 1.7
        kappa = torch.exp(torch.rand(B, generator=g) * 4 + 1.5)  # ~ [4.5, 245]
        W = torch.randn(N, B, d, generator=g)
        locp = loc.clone().requires_grad_(True)
        torch.manual_seed(100)
        dist = VonMisesFisher(locp, kappa)
        s = dist.rsample((N,))
        (s * W).sum().backward()
        v, e, u = replay_vmf(100, B, N, d)
        kp = kappa.clone().requires_grad_(True)
        dist2 = VonMisesFisher(loc, kp)
        ent = dist2.entropy()
        ent.sum().backward()
        lp = VonMisesFisher(loc, kappa).log_prob(s.detach()[0])
        out[tag] = dict(loc=loc, kappa=kappa, W=W, samples=s.detach(), gloc=locp.grad, v=v, eps=e,
                        u=u, entropy=ent.detach(), gkappa=kp.grad, log_prob=lp)

    # ---- log-partition at d = 512 (nu = 255) over the model's kappa range ---------------------
    kap = torch.tensor([10.0, 37.5, 200.0, 1000.0, 2.0e3, 1.0e5], requires_grad=True)
    logC = vMFLogPartition.apply(512, kap)
    logC.sum().backward()
    out["logc_d512"] = dict(kappa=kap.detach(), logC=logC.detach(), dlogC=kap.grad)
    kap2 = torch.tensor([0.05, 1.0, 8.0, 50.0, 400.0], requires_grad=True)
    logC2 = vMFLogPartition.apply(3, kap2)
    logC2.sum().backward()
    out["logc_d3"] = dict(kappa=kap2.detach(), logC=logC2.detach(), dlogC=kap2.grad)

    # ---- ProjectedNormal ---------------------------------------------------------------------
    B, d, N = 6, 24, 4
    mu = torch.randn(B, d, generator=g)
    sigma = torch.exp(torch.randn(B, d, generator=g) * 0.5 - 2.0)
    W = torch.randn(N, B, d, generator=g)
    mup, sgp = mu.clone().requires_grad_(True), sigma.clone().requires_grad_(True)
    torch.manual_seed(200)
    pn = ProjectedNormal(mup, sgp)
    s = pn.rsample((N,))
    (s * W).sum().backward()
    eps = replay_pn(200, (N, B, d))
    out["pn"] = dict(mu=mu, sigma=sigma, W=W, samples=s.detach(), gmu=mup.grad, gsigma=sgp.grad,
                     eps=eps, entropy=ProjectedNormal(mu, sigma).entropy(),
                     loc=ProjectedNormal(mu, sigma).loc)

    # ---- VClipLoss end to end (vMF, ProjectedNormal) -----------------------------------------
    B, d, S = 8, 16, 3
    for eps_ls in (0.1, 0.0):
        li = torch.randn(B, d, generator=g)
        lt = li + 0.4 * torch.randn(B, d, generator=g)
        ki = torch.exp(torch.rand(B, generator=g) * 3 + 2.0)
        kt = torch.exp(torch.rand(B, generator=g) * 3 + 2.0)
        ls = torch.tensor(2.6592)
        leaves = [t.clone().requires_grad_(True) for t in (li, lt, ki, kt, ls)]
        crit = VClipLoss(num_samples=S, label_smoothing=eps_ls)
        torch.manual_seed(300)
        res = crit(VonMisesFisher(leaves[0], leaves[2]), VonMisesFisher(leaves[1], leaves[3]),
                   1.0 / leaves[2], 1.0 / leaves[3], leaves[4], True)
        res["total_loss"].backward()
        torch.manual_seed(300)  # image rsample then text rsample
        vi, ei, ui = [], [], []
        noise = []
        for _ in range(2):
            vs, es, us = [], [], []
            for _b in range(B):
                vs.append(torch.randn(S, d - 1))
                beta = torch.distributions.Beta(torch.tensor(0.5 * (d - 1)),
                                                torch.tensor(0.5 * (d - 1)))
                uni = torch.distributions.Uniform(torch.tensor(0.0), torch.tensor(1.0))
                es.append(beta.sample([10 * S]))
                us.append(uni.sample([10 * S]))
            noise.append((torch.stack(vs), torch.stack(es), torch.stack(us)))
        out["vclip_vmf_eps%g" % eps_ls] = dict(
            li=li, lt=lt, ki=ki, kt=kt, ls=ls, vi=noise[0][0], ei=noise[0][1], ui=noise[0][2],
            vt=noise[1][0], et=noise[1][1], ut=noise[1][2],
            **{k: v.detach() for k, v in res.items()},
            **{"g" + n: l.grad for n, l in zip(("li", "lt", "ki", "kt", "ls"), leaves)})

    mi = torch.randn(B, d, generator=g)
    mt = mi + 0.3 * torch.randn(B, d, generator=g)
    si = torch.exp(torch.randn(B, d, generator=g) * 0.3 - 2.0)
    st = torch.exp(torch.randn(B, d, generator=g) * 0.3 - 2.0)
    ls = torch.tensor(3.0)
    leaves = [t.clone().requires_grad_(True) for t in (mi, mt, si, st, ls)]
    crit = VClipLoss(num_samples=S, label_smoothing=0.1)
    torch.manual_seed(400)
    res = crit(ProjectedNormal(leaves[0], leaves[2]), ProjectedNormal(leaves[1], leaves[3]),
               leaves[2].mean(-1), leaves[3].mean(-1), leaves[4], True)
    res["total_loss"].backward()
    torch.manual_seed(400)
    e_i = torch.distributions.utils._standard_normal((S, B, d), dtype=torch.float32, device="cpu")
    e_t = torch.distributions.utils._standard_normal((S, B, d), dtype=torch.float32, device="cpu")
    out["vclip_pn"] = dict(mi=mi, mt=mt, si=si, st=st, ls=ls, ei=e_i, et=e_t,
                           **{k: v.detach() for k, v in res.items()},
                           **{"g" + n: l.grad for n, l in zip(("mi", "mt", "si", "st", "ls"),
                                                              leaves)})
    # eval branch (is_train=False): means only
    res = VClipLoss(label_smoothing=0.1)(VonMisesFisher(li, ki), VonMisesFisher(lt, kt), 1.0 / ki,
                                         1.0 / kt, torch.tensor(2.6592), False)
    out["vclip_vmf_eval"] = dict(li=li, lt=lt, ki=ki, kt=kt,
                                 **{k: v.detach() for k, v in res.items()})

    flat = {}
    for t, c in out.items():
        for k, v in c.items():
            flat[f"{t}/{k}"] = v.detach().numpy() if torch.is_tensor(v) else np.asarray(v)
    np.savez_compressed(os.path.join(OUT, "dist.npz"), **flat)
    print("wrote dist.npz with", len(flat), "arrays")


if __name__ == "__main__":
    main()

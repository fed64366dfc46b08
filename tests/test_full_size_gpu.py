"""GPU, BASELINE.json full size (CLIP B=32768 d=512 bf16): the N x N oracle does not fit in
seconds, so parity is checked through size-independent properties and slab oracles:
  * row / column LSE and positives of a random slab of rows / columns against a direct fp32
    evaluation of that slab;
  * d loss / d logit_scale against a central finite difference of two extra forwards;
  * gradients of a slab of rows against the oracle formula G = P - target evaluated on the slab
    from the full statistics;
  * permutation equivariance: permuting the pairs leaves the loss unchanged and permutes grads.
"""
import pytest
import torch

from oracle import losses_oracle as O

pytestmark = pytest.mark.gpu

N, D, LS, EPS = 32768, 512, 2.6592, 0.1


@pytest.fixture(scope="module")
def full(lib_built):
    from multimodal_2025_b200.losses import ClipLoss
    x, y = O.synth_embeddings(N, D, 1234, correlated=True, dtype=torch.bfloat16)
    xc = x.cuda().requires_grad_(True)
    yc = y.cuda().requires_grad_(True)
    ls = torch.tensor(LS, device="cuda", requires_grad=True)
    crit = ClipLoss(EPS).cuda()
    loss = crit(xc, yc, ls)
    loss.backward()
    torch.cuda.synchronize()
    return dict(crit=crit, x=xc, y=yc, ls=ls, loss=loss.detach())


def _normalized(t):
    return torch.nn.functional.normalize(t.detach().float(), dim=-1).bfloat16().float()


def test_loss_matches_blockwise_fp32_oracle(full):
    """the reference formula (clipLoss.py:106-119) evaluated block by block in fp32 on the GPU."""
    xh, yh = _normalized(full["x"]), _normalized(full["y"])
    s = float(torch.tensor(LS).exp())
    r = torch.empty(N, device="cuda")
    cm = torch.full((N,), -float("inf"), device="cuda")
    cs = torch.zeros(N, device="cuda")
    tot = torch.zeros((), device="cuda", dtype=torch.float64)
    diag = torch.empty(N, device="cuda")
    for i0 in range(0, N, 4096):
        z = s * (xh[i0:i0 + 4096] @ yh.T)
        r[i0:i0 + 4096] = torch.logsumexp(z, dim=1)
        m = torch.maximum(cm, z.max(dim=0).values)
        cs = cs * (cm - m).exp() + (z - m).exp().sum(dim=0)
        cm = m
        tot += z.double().sum()
        diag[i0:i0 + 4096] = z[torch.arange(4096), torch.arange(i0, i0 + 4096)]
    c = cm + cs.log()
    ref = ((r.double().sum() + c.double().sum()) / (2 * N) - (1 - EPS) / N * diag.double().sum()
           - EPS / N ** 2 * tot)
    assert abs(float(full["loss"]) - float(ref)) / abs(float(ref)) < 2e-3


def test_logit_scale_gradient_vs_finite_difference(full):
    crit, x, y = full["crit"], full["x"].detach(), full["y"].detach()
    h = 1e-2
    with torch.no_grad():
        lp = crit(x, y, torch.tensor(LS + h, device="cuda")).double()
        lm = crit(x, y, torch.tensor(LS - h, device="cuda")).double()
    fd = float((lp - lm) / (2 * h))
    assert abs(float(full["ls"].grad) - fd) <= 2e-2 * abs(fd)


def test_gradient_slab_vs_oracle_formula(full):
    """rows [5000, 5256): dL/dxhat = s G Yhat with G from the closed form,
    pushed through the normalisation; compares with the kernel's dL/dx for those rows."""
    xh, yh = _normalized(full["x"]), _normalized(full["y"])
    s = float(torch.tensor(LS).exp())
    rows = slice(5000, 5256)
    # column LSE over ALL rows (blockwise), row LSE for the slab
    cm = torch.full((N,), -float("inf"), device="cuda")
    cs = torch.zeros(N, device="cuda")
    for i0 in range(0, N, 4096):
        z = s * (xh[i0:i0 + 4096] @ yh.T)
        m = torch.maximum(cm, z.max(dim=0).values)
        cs = cs * (cm - m).exp() + (z - m).exp().sum(dim=0)
        cm = m
    c = cm + cs.log()
    z = s * (xh[rows] @ yh.T)
    r = torch.logsumexp(z, dim=1, keepdim=True)
    G = (0.5 / N) * ((z - r).exp() + (z - c[None, :]).exp()) - EPS / N ** 2
    idx = torch.arange(5000, 5256, device="cuda")
    G[torch.arange(256, device="cuda"), idx] -= (1 - EPS) / N
    gx_hat = s * (G @ yh)
    xr = full["x"].detach().float()[rows]
    nrm = xr.norm(dim=-1, keepdim=True)
    xhr = xr / nrm
    ref = (gx_hat - xhr * (xhr * gx_hat).sum(-1, keepdim=True)) / nrm
    got = full["x"].grad[rows].float()
    assert float((got - ref).norm() / ref.norm()) < 2e-2


def test_permutation_equivariance(full):
    crit = full["crit"]
    perm = torch.randperm(N, generator=torch.Generator().manual_seed(0)).cuda()
    xp = full["x"].detach()[perm].requires_grad_(True)
    yp = full["y"].detach()[perm].requires_grad_(True)
    ls = torch.tensor(LS, device="cuda")
    lp = crit(xp, yp, ls)
    lp.backward()
    assert abs(float(lp) - float(full["loss"])) < 1e-4 * abs(float(full["loss"]))
    ref = full["x"].grad[perm].float()
    assert float((xp.grad.float() - ref).norm() / ref.norm()) < 1e-2

"""Diagnostic: gradient error of the kept-E backward vs the fp32 closed form for several shapes,
per output (da, db), and where the error sits.  python tools/diag_te.py"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from multimodal_2025_b200 import ops

def run(B, M, N, d, off, ls):
    g = torch.Generator().manual_seed(7)
    a = torch.nn.functional.normalize(torch.randn(B, M, d, generator=g), dim=-1).cuda().bfloat16()
    b = torch.nn.functional.normalize(torch.randn(B, N, d, generator=g), dim=-1).cuda().bfloat16()
    scale = torch.tensor([ls], device="cuda")
    bound = ops.DOT_BOUND_NORMALIZED
    row_lse, cmax, csum, diag, e_keep, e_flag = ops.sim_fwd_keep(
        a, b, scale, ops.SCALE_LOG, 4.6052, off, True, ops.IMPL_TC, bound)
    col_lse = cmax + csum.log()
    nn_ = float(N)
    w = (0.5 / nn_, 0.5 / nn_, 0.9 / nn_, 0.1 / (nn_ * nn_))
    gout = torch.linspace(0.5, 1.5, B, device="cuda")
    ref = ops.sim_bwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, off,
                      row_lse, col_lse, *w, gout, True, True, False, ops.IMPL_TC, bound)
    got = ops.sim_bwd(a, b, None, None, None, scale, ops.SCALE_LOG, 4.6052, ops.SIM_PLAIN, off,
                      row_lse, col_lse, *w, gout, True, True, False, ops.IMPL_TC, bound,
                      e_keep, e_flag)
    s = float(torch.tensor(ls).clamp(max=4.6052).exp())
    af, bf = a.float(), b.float()
    z = s * (af @ bf.transpose(1, 2))
    r = torch.logsumexp(z, dim=2, keepdim=True)
    c = torch.logsumexp(z, dim=1, keepdim=True)
    G = w[0] * (z - r).exp() + w[1] * (z - c).exp() - w[3]
    idx = torch.arange(M, device="cuda")
    G[:, idx, idx + off] -= w[2]
    G = G * gout.reshape(B, 1, 1) * s
    exact = (G @ bf, G.transpose(1, 2) @ af)
    out = []
    for k in (0, 1):
        den = exact[k].norm().item()
        e_rec = (ref[k] - exact[k]).norm().item() / den
        e_kept = (got[k] - exact[k]).norm().item() / den
        # per 256-row block error of the kept path
        diff = (got[k] - exact[k]).reshape(-1, got[k].shape[-1])
        ex = exact[k].reshape(-1, got[k].shape[-1])
        blk = [(diff[i:i + 256].norm() / ex[i:i + 256].norm()).item() for i in range(0, diff.shape[0], 256)]
        cblk = [(diff[:, i:i + 256].norm() / ex[:, i:i + 256].norm()).item() for i in range(0, diff.shape[1], 256)]
        out.append((round(e_rec, 5), round(e_kept, 5), [round(x, 4) for x in blk[:6]], [round(x, 4) for x in cblk]))
    print((B, M, N, d, off, ls), "flag", int(e_flag.item()), "dscale", float(ref[2]), float(got[2]), out)

for cfg in [(1, 256, 256, 256, 0, 2.6592), (1, 512, 1024, 512, 256, 2.6592), (2, 512, 512, 1024, 0, 3.0),
            (1, 768, 1024, 256, 0, 2.0), (1, 768, 1280, 768, 0, 2.6592), (1, 4096, 4096, 512, 0, 2.6592),
            (1, 2048, 2048, 512, 0, 4.0)]:
    run(*cfg)

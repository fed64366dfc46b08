#!/usr/bin/env python
"""Secondary measurements of the other SURVEY §8 rows (one B200): sampler kernels against the HBM
roofline (algorithmic bytes), vCLIP / CLOOB / AlignCLIP / top-k fwd+bwd times at the
BASELINE configs (or the largest size the fp32 SIMT path finishes in seconds, stated per line).
Prints one JSON object per line; `python tools/bench_paths.py > profiles/rN_paths.jsonl`."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import multimodal_2025_b200  # noqa: E402,F401
from multimodal_2025_b200 import ops  # noqa: E402
from multimodal_2025_b200.losses import ClipLoss, CLOOBLoss  # noqa: E402
from multimodal_2025_b200.losses.align_clip_loss import ClipInModalityLoss  # noqa: E402
from multimodal_2025_b200.losses.distributions import (PowerSpherical, ProjectedNormal,  # noqa: E402
                                                       VonMisesFisher)
from multimodal_2025_b200.losses.vclipLoss import VClipLoss  # noqa: E402

PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps


def emit(**kw):
    print(json.dumps(kw), flush=True)


def hbm_line(name, ms, algo_bytes, note):
    gbs = algo_bytes / (ms * 1e-3) / 1e9
    emit(path=name, ms=ms, roofline={"bound": "hbm", "achieved": gbs, "peak": PEAKS["hbm_gbs"],
                                     "unit": "GB/s", "frac": gbs / PEAKS["hbm_gbs"]},
         algorithmic_bytes=algo_bytes, note=note)


def main():
    g = torch.Generator().manual_seed(0)
    S, B, d = 20, 8192, 512  # cfg4
    loc = torch.nn.functional.normalize(torch.randn(B, d, generator=g), dim=-1).to(dev)
    kappa = torch.exp(torch.rand(B, generator=g) * 5.3 + 2.3).to(dev)  # LogUniform[10, 2000]
    sigma = torch.exp(torch.rand(B, d, generator=g) * 3.4 - 4.6).to(dev)
    gout = torch.randn(S, B, d, generator=g).to(dev)
    out_b, par_b = S * B * d * 4, B * d * 4

    # ---- sampler kernels alone (C-ABI ops), Philox noise -------------------------------------
    ms = timeit(lambda: ops.projnormal_fwd(loc, sigma, None, 1, S))
    hbm_line("sample_projnormal_fwd", ms, out_b + 2 * par_b, "cfg4 S=20 B=8192 d=512 fp32")
    y, inv = ops.projnormal_fwd(loc, sigma, None, 1, S)
    ms = timeit(lambda: ops.projnormal_bwd(gout, y, inv, None, 1))
    hbm_line("sample_projnormal_bwd", ms, 2 * out_b + 2 * par_b, "reads g, y; writes dmu, dsigma")
    t = torch.rand(S, B, device=dev) * 1.6 - 0.8
    ms = timeit(lambda: ops.sphere_fwd(t, loc, None, 1, *ops.POWER_SPHERICAL_CFG))
    hbm_line("sphere_sample_fwd", ms, out_b + par_b, "PowerSpherical / vMF transform")
    ys = ops.sphere_fwd(t, loc, None, 1, *ops.POWER_SPHERICAL_CFG)
    ms = timeit(lambda: ops.sphere_bwd(gout, ys, t, loc, *ops.POWER_SPHERICAL_CFG))
    hbm_line("sphere_sample_bwd", ms, 2 * out_b + 2 * par_b, "reads g, y; writes dloc, dt")
    ms = timeit(lambda: ops.vmf_sample_w(kappa, d, S, 1, None, None))
    emit(path="vmf_sample_w (Wood rejection, Philox)", ms=ms, note="S*B = 163840 scalars")
    a = (d - 1) / 2 + kappa
    b = torch.full_like(kappa, (d - 1) / 2)
    ms = timeit(lambda: ops.beta_sample(a, b, S, 1))
    emit(path="beta_sample", ms=ms, note="S*B scalars")
    z = ops.beta_sample(a, b, S, 1)
    ms = timeit(lambda: ops.beta_grad(z, a, b))
    emit(path="beta_grad", ms=ms, note="S*B scalars")
    ms = timeit(lambda: ops.log_bessel(kappa, d / 2 - 1))
    emit(path="log_bessel (nu=255, B=8192)", ms=ms, note="fp64 Debye expansion, both orders")

    # ---- distribution rsample fwd+bwd through the drop-in classes ----------------------------
    for name, mk in (("PowerSpherical", lambda l, k: PowerSpherical(l, k)),
                     ("VonMisesFisher", lambda l, k: VonMisesFisher(l, k)),
                     ("ProjectedNormal", lambda l, k: ProjectedNormal(l, sigma))):
        def f():
            l = loc.clone().requires_grad_(True)
            k = kappa.clone().requires_grad_(True)
            smp = mk(l, k).rsample((S,))
            smp.backward(gout)
        ms = timeit(f, reps=5)
        emit(path="%s.rsample((20,)) fwd+bwd" % name, ms=ms,
             note="cfg4 B=8192 d=512; reference vMF is a Python loop at ~1.2 ms/row")

    # ---- vCLIP end to end (cfg4), fp32 samples -> bf16 tensor-core CLIP kernels ---------------
    for name, bf16 in (("power_spherical", True), ("vmf", True), ("projected_normal", True),
                       ("power_spherical", False)):
        crit = VClipLoss(num_samples=S, label_smoothing=0.0).to(dev)
        if bf16:
            crit.clip_loss.compute_dtype = torch.bfloat16
        ls = torch.tensor(2.6592, device=dev, requires_grad=True)

        def f():
            li = loc.clone().requires_grad_(True)
            lt = loc.clone().requires_grad_(True)
            k = kappa.clone().requires_grad_(True)
            if name == "power_spherical":
                di, dt_ = PowerSpherical(li, k), PowerSpherical(lt, k)
            elif name == "vmf":
                di, dt_ = VonMisesFisher(li, k), VonMisesFisher(lt, k)
            else:
                di, dt_ = ProjectedNormal(li, sigma), ProjectedNormal(lt, sigma)
            v = 1.0 / k if name != "projected_normal" else sigma.mean(-1)
            crit(di, dt_, v, v, ls, True)["total_loss"].backward()
        try:
            ms = timeit(f, reps=2, warm=1)
            flops = S * 6.0 * B * B * d
            emit(path="VClipLoss[%s] fwd+bwd" % name, ms=ms, pairs_per_s=B / (ms * 1e-3),
                 tflops_algorithmic=flops / (ms * 1e-3) / 1e12,
                 note="cfg4 S=20 B=8192 d=512; fp32 samples; CLIP term on " +
                      ("bf16 tcgen05 kernels (compute_dtype=bf16, 2e-3 contract)" if bf16 else
                       "fp32 SIMT kernels (reference dtype, 1e-5 contract)"))
        except Exception as e:  # noqa: BLE001
            emit(path="VClipLoss[%s]" % name, error=repr(e)[:200])

    # ---- CLOOB (cfg3) ------------------------------------------------------------------------
    n = 4096
    x = torch.nn.functional.normalize(torch.randn(n, d, generator=g), dim=-1).to(dev)
    yv = torch.nn.functional.normalize(torch.randn(n, d, generator=g), dim=-1).to(dev)
    crit = CLOOBLoss(30.0, 15.0, dev)

    def f():
        a_, b_ = x.clone().requires_grad_(True), yv.clone().requires_grad_(True)
        crit(a_, b_).backward()
    ms = timeit(f, reps=3, warm=1)
    emit(path="CLOOBLoss fwd+bwd", ms=ms, pairs_per_s=n / (ms * 1e-3),
         tflops_algorithmic=60.0 * n * n * d / (ms * 1e-3) / 1e12,
         note="cfg3 N=4096 d=512 fp32 SIMT path (60 N^2 d algorithmic FLOP)")
    xb, yb = x.bfloat16(), yv.bfloat16()

    def fb():
        a_, b_ = xb.clone().requires_grad_(True), yb.clone().requires_grad_(True)
        crit(a_, b_).backward()
    ms = timeit(fb, reps=5, warm=2)
    emit(path="CLOOBLoss fwd+bwd", ms=ms, pairs_per_s=n / (ms * 1e-3),
         tflops_algorithmic=60.0 * n * n * d / (ms * 1e-3) / 1e12,
         note="cfg3 N=4096 d=512 bf16: Hopfield retrievals + InfoLOOB on the tcgen05 kernels "
              "(P-panel GEMM forward, dual-accumulator sweep backward, MASKDIAG)")

    from multimodal_2025_b200.graphs import GraphedStep  # noqa: E402
    gs = GraphedStep(lambda a_, b_: crit(a_, b_), [xb.clone().requires_grad_(True),
                                                    yb.clone().requires_grad_(True)])
    ms = timeit(lambda: gs(xb, yb), reps=5, warm=2)
    emit(path="CLOOBLoss fwd+bwd (CUDA-graph replay)", ms=ms, pairs_per_s=n / (ms * 1e-3),
         tflops_algorithmic=60.0 * n * n * d / (ms * 1e-3) / 1e12,
         note="cfg3 N=4096 d=512 bf16, graphs.GraphedStep")

    # ---- AlignCLIP semantic (reduced N: fp32 SIMT MODULATED kernels) ---------------------------
    n, da_, dsem = 8192, 1024, 768
    xi = torch.nn.functional.normalize(torch.randn(n, da_, generator=g), dim=-1).to(dev)
    xt = torch.nn.functional.normalize(torch.randn(n, da_, generator=g), dim=-1).to(dev)
    sem = torch.randn(n, dsem, generator=g).to(dev)
    crit = ClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=True)
    sc = torch.tensor(14.285, device=dev)

    def f():
        a_, b_ = xi.bfloat16().requires_grad_(True), xt.bfloat16().requires_grad_(True)
        crit(a_, b_, sc, semantic_features=sem).backward()
    ms = timeit(f, reps=3, warm=1)
    emit(path="ClipInModalityLoss(semantic) fwd+bwd", ms=ms, pairs_per_s=n / (ms * 1e-3),
         tflops_algorithmic=n * n * (12.0 * da_ + 2 * dsem) / (ms * 1e-3) / 1e12,
         note="N=8192 d=1024 d_sem=768 bf16 (cfg5 is N=65536 over 8 GPUs); cross-modal term on "
              "the PLAIN tcgen05 kernels, in-modality MODULATED term on the two-accumulator "
              "tcgen05 kernels (csrc/sim_tc_mod.cu)")

    xa, xtb = xi.bfloat16(), xt.bfloat16()
    gs = GraphedStep(lambda a_, b_: crit(a_, b_, sc, semantic_features=sem),
                     [xa.clone().requires_grad_(True), xtb.clone().requires_grad_(True)])
    ms = timeit(lambda: gs(xa, xtb), reps=5, warm=2)
    emit(path="ClipInModalityLoss(semantic) fwd+bwd (CUDA-graph replay)", ms=ms,
         pairs_per_s=n / (ms * 1e-3),
         tflops_algorithmic=n * n * (12.0 * da_ + 2 * dsem) / (ms * 1e-3) / 1e12,
         note="N=8192 d=1024 d_sem=768 bf16, graphs.GraphedStep")

    # ---- top-k ---------------------------------------------------------------------------------
    Q, Nt, k = 1024, 65536, 100
    scores = torch.randn(Q, Nt, generator=g).to(dev)
    ms = timeit(lambda: ops.topk_rows(scores, k))
    hbm_line("topk_rows k=100", ms, Q * Nt * 4, "Q=1024 N_t=65536 fp32 scores read once")
    ms_t = timeit(lambda: torch.topk(scores, k, dim=1))
    emit(path="torch.topk same input (library baseline)", ms=ms_t)

    # ---- retrieval ranks without the sort (rank.cu) ----------------------------------------------
    del scores
    from multimodal_2025_b200.metrics import get_clip_metrics, retrieval_ranks
    for n in (8192, 32768):
        sc = torch.randn(n, n, generator=g).to(dev)
        ar = torch.arange(n, device=dev)
        ms = timeit(lambda: retrieval_ranks(sc, ar, ar), reps=5, warm=2)
        hbm_line("rank_both (rows + columns) N=%d" % n, ms, n * n * 4,
                 "fp32 scores read once; includes the two target gathers and two memsets")
        if n == 8192:  # what the reference does with the same matrix (train.py:348-350), on GPU
            ms_s = timeit(lambda: (torch.argsort(sc, descending=True),
                                   torch.argsort(sc.t(), descending=True)), reps=3, warm=1)
            emit(path="2 x torch.argsort of the same N=8192 matrix (library baseline)", ms=ms_s)
        del sc
    n, dd = 32768, 512
    xi = torch.nn.functional.normalize(torch.randn(n, dd, generator=g), dim=-1).to(dev)
    xt = torch.nn.functional.normalize(xi + 0.05 * torch.randn(n, dd, device=dev), dim=-1)
    ms = timeit(lambda: get_clip_metrics(xi, xt, 14.2857), reps=3, warm=1)
    emit(path="get_clip_metrics N=32768 d=512 fp32 (GEMM + rank_both + 2 x D2H of N ints)", ms=ms)
    # fused GEMM + rank count (csrc/sim_tc_rank.cu): no score matrix; one sweep per direction
    from multimodal_2025_b200.metrics.retrieval import fused_ranks
    xb, tb = xi.bfloat16(), xt.bfloat16()
    ar = torch.arange(n, device=dev)
    ms = timeit(lambda: fused_ranks(xb, tb, ar), reps=5, warm=2)
    tf = 2.0 * n * n * dd / (ms * 1e-3) / 1e12
    emit(path="fused_ranks (sim_rank_tc_kernel, one direction) N=32768 d=512 bf16", ms=ms,
         roofline={"bound": "tensor", "achieved": tf, "peak": PEAKS["bf16_tflops"],
                   "unit": "TFLOP/s", "frac": tf / PEAKS["bf16_tflops"]},
         note="includes the fp32 target dot products and the memset; scores never written")
    ms = timeit(lambda: get_clip_metrics(xb, tb, 14.2857, fused=True), reps=3, warm=1)
    emit(path="get_clip_metrics N=32768 d=512 bf16 FUSED (2 sweeps + 2 x D2H of N ints)", ms=ms)
    # small-Q top-k (the reference's call shape: one query against all texts)
    for Q in (1, 16):
        sc1 = torch.randn(Q, 1 << 20, generator=g).to(dev)
        ms = timeit(lambda: ops.topk_rows(sc1, 5))
        hbm_line("topk_rows k=5 Q=%d N_t=1048576" % Q, ms, Q * (1 << 20) * 4, "small-Q shape")


if __name__ == "__main__":
    main()

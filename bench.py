#!/usr/bin/env python
"""bench.py — contrastive-loss fwd+bwd pairs/s on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--config cfg1|cfg2|cfg3|cfg4|cfg5] [--precision ref|bf16] [--vdist NAME]

Workloads (BASELINE.json `configs`); the default, cfg2, is the one the metric is
quoted on:
  cfg1  CLIP  N=256   d=512  fp32          (the reference's own CPU-runnable case)
  cfg2  CLIP  N=32768 d=512  bf16          (headline; row-sharded over the ranks at --gpus N)
  cfg3  CLOOB N=4096  d=512  fp32 inputs   (InfoLOOB + 4 Hopfield retrievals)
  cfg4  vCLIP B=8192  d=512  S=20 samples  (--vdist power_spherical|vmf|projected_normal)
  cfg5  AlignCLIP N=65536 d=1024 d_sem=768 bf16 (row-sharded over the ranks)
cfg3 / cfg4 run at the REFERENCE's precision (fp32 inputs -> fp32-grade arithmetic: the CLIP /
vCLIP similarity kernels on the tensor cores over two-term bf16 splits, CLOOB on the fp32 FFMA
kernels) unless `--precision bf16`, which feeds the tensor-core kernels plain bf16 operands and
is labelled as reduced precision.

A "step" is one forward+backward of the loss over one synthetic batch.  `value` is timed with
CUDA events per step, inputs resident in HBM, an L2 flush (256 MB write) before every step; `e2e`
goes through the public loss module with HOST (pinned) inputs, the H2D copies and the D2H read
of the loss inside the timed region.  `roofline` is the dominant kernel of the step, timed INSIDE
eager steps by the library's per-launch CUDA events (mm25_debug_timing, on the launch stream).
`--impl reference` times the CPU restatement of the reference (oracle/, pinned by the goldens
the reference itself produced) on the host cores, on a bounded sample of the same workload.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "CLIP-loss fwd+bwd pairs/s at B=32k d=512, 1/2/4/8 B200; % of bf16 TC peak"
LOGIT_SCALE = 2.6592  # log(1/0.07), CLIP init

CONFIGS = {
    "cfg1": dict(loss="clip", n=256, d=512, dtype="f32", scaling="strong"),
    "cfg2": dict(loss="clip", n=32768, d=512, dtype="bf16", scaling="strong"),
    "cfg3": dict(loss="cloob", n=4096, d=512, dtype="f32", scaling="replicas"),
    "cfg4": dict(loss="vclip", n=8192, d=512, S=20, dtype="f32", scaling="replicas"),
    "cfg5": dict(loss="alignclip", n=65536, d=1024, dsem=768, dtype="bf16", scaling="strong"),
}
SLOT_NAMES = ["sim_fwd_tc_kernel", "gemm_te_kernel", "sim_grad_tc_kernel", "gemm_grouped_kernel",
              "sim_mod_tc_kernel", "sim_fwd_kernel (fp32 FFMA)", "sim_bwd_kernel (fp32 FFMA)",
              "sampler kernels"]


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    p.add_argument("--precision", default="ref", choices=["ref", "bf16"],
                   help="cfg3/cfg4: ref = the reference's fp32 arithmetic; bf16 = tensor-core "
                        "operands in bf16 (labelled reduced precision)")
    p.add_argument("--vdist", default="power_spherical",
                   choices=["power_spherical", "vmf", "projected_normal"])
    # --batch / --dim: `--n`, `--d` are kept for plain runs, but torchrun's own parser rejects them
    # as ambiguous abbreviations of its options
    p.add_argument("--n", "--batch", dest="n", type=int, default=0,
                   help="override the batch (pairs)")
    p.add_argument("--d", "--dim", dest="d", type=int, default=0)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-graph", action="store_true", help="eager launches instead of a CUDA graph")
    p.add_argument("--serial-e2e", action="store_true",
                   help="e2e without the double-buffered input pipeline (H2D, step, D2H in series)")
    p.add_argument("--slab-rows", type=int, default=2048, help="rows of the CPU sample slab")
    a = p.parse_args()
    a.cfg = dict(CONFIGS[a.config])
    if a.n:
        a.cfg["n"] = a.n
    if a.d:
        a.cfg["d"] = a.d
    return a


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            pk = json.load(f)
        return {"bf16_tflops": pk.get("bf16_tflops", 1590.0),
                "bf16_tflops_sustained": pk.get("bf16_tflops_sustained", 1400.0),
                "hbm_gbs": pk.get("hbm_gbs", 6650.0), "source": "measured"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md)"}


def synth_pair(n, d, seed):
    """seeded L2-normalised image / text embeddings (fp32, CPU generator)."""
    import torch
    import torch.nn.functional as F
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = F.normalize(torch.randn(n, d, generator=g, dtype=torch.float32), dim=-1)
    y = F.normalize(torch.randn(n, d, generator=g, dtype=torch.float32), dim=-1)
    return x, y


def workload_name(cfg, args):
    c = cfg
    if c["loss"] == "clip":
        return ("CLIP InfoNCE fwd+bwd, global B=%d d=%d %s, label_smoothing 0.1, logit_scale "
                "log(1/0.07)" % (c["n"], c["d"], c["dtype"]))
    if c["loss"] == "cloob":
        return ("CLOOB (InfoLOOB + Hopfield retrieval) fwd+bwd, B=%d d=%d, inv_tau 30, "
                "scale_hopfield 15, %s" % (c["n"], c["d"], precision_note(c, args)))
    if c["loss"] == "vclip":
        return ("vCLIP fwd+bwd, %s sampling S=%d, B=%d d=%d, %s"
                % (args.vdist, c["S"], c["n"], c["d"], precision_note(c, args)))
    return ("AlignCLIP (cross-modal CE + semantic in-modality CE) fwd+bwd, global B=%d d=%d "
            "d_sem=%d bf16" % (c["n"], c["d"], c["dsem"]))


def precision_note(c, args):
    if c["dtype"] == "f32" and args.precision == "bf16":
        return "fp32 inputs, similarity GEMMs on bf16 tensor-core operands (REDUCED precision)"
    if c["dtype"] == "f32":
        if c["loss"] in ("clip", "vclip") and os.environ.get("MM25_FP32_TC", "1") != "0":
            return ("fp32 inputs, fp32-grade arithmetic (reference precision: similarity GEMMs "
                    "on the tensor cores over two-term bf16 splits, 16 significant bits, fp32 "
                    "accumulation; 1e-5 / 1e-4 contract)")
        return "fp32 inputs and arithmetic (reference precision, fp32 FFMA kernels)"
    return "bf16"


def algorithmic_flops(c):
    n, d = float(c["n"]), float(c["d"])
    if c["loss"] == "clip":
        return 6.0 * n * n * d
    if c["loss"] == "cloob":
        return 60.0 * n * n * d
    if c["loss"] == "vclip":
        return c["S"] * 6.0 * n * n * d
    return n * n * (12.0 * d + 2.0 * c["dsem"])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
                pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "sm_mhz_min": sm[0] if sm else None,
                "power_w_max": max(pw) if pw else None}


# ------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the pinned oracle (kind "port") on the host cores
# ------------------------------------------------------------------------------------------
def _best(fn, reps):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return ts


def cpu_clip_full(n, d, seed=1234):
    """fwd+bwd of oracle.losses_oracle.clip_loss (reference clipLoss.py:81-121) on the whole batch."""
    import torch
    from oracle import losses_oracle as O
    x, y = synth_pair(n, d, seed)
    ls = torch.tensor(LOGIT_SCALE)

    def step():
        O.loss_and_grads(O.clip_loss, [x, y, ls], 0.1, True)
    return step


def cpu_clip_slab(n, d, rows, seed=1234):
    """fwd+bwd of oracle.losses_oracle.clip_loss_slab: the [rows, N] slab of the same logits with
    the same per-logit arithmetic (rows/N of one step's N x N work; the slabs recombine to the
    exact clip_loss, tests/test_oracle.py)."""
    import torch
    from oracle import losses_oracle as O
    x, y = synth_pair(n, d, seed)
    ls = torch.tensor(LOGIT_SCALE)

    def step():
        xs, ys = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
        l = ls.clone().requires_grad_(True)
        img_sum, col_lse, col_lin, diag = O.clip_loss_slab(xs, ys, l, 0, rows, 0.1, True)
        # the slab's share of the full loss: image-side CE of its rows + its column terms
        part = 0.5 * (img_sum / n + (col_lse.mean() - 0.9 * diag.sum() / n - 0.1 / n * col_lin.mean()))
        part.backward()
    return step


def cpu_case(args):
    """Returns (step_fn, pairs_per_step_equivalent, sample_description) for --impl reference and
    the cpu_baseline object: a bounded sample of the selected workload through the oracle."""
    import torch
    c = args.cfg
    n, d = c["n"], c["d"]
    if c["loss"] == "clip":
        if n <= 8192:
            return cpu_clip_full(n, d), float(n), "whole batch: oracle.clip_loss fwd+bwd, N=%d fp32" % n
        rows = min(args.slab_rows, n)
        return (cpu_clip_slab(n, d, rows), float(rows),
                "slab of %d of %d image rows x all %d texts (%.4f of one step's N x N work, same "
                "per-pair arithmetic; oracle.clip_loss_slab fwd+bwd, fp32)" % (rows, n, n, rows / n))
    from oracle import losses_oracle as O
    if c["loss"] == "cloob":
        x, y = synth_pair(n, d, 1234)

        def step():
            O.loss_and_grads(lambda a, b: O.cloob_loss(a, b, 30.0, 15.0), [x, y])
        return step, float(n), "whole batch: oracle.cloob_loss fwd+bwd, N=%d fp32" % n
    if c["loss"] == "vclip":
        from oracle import dist_oracle as DO
        S, Ss = c["S"], 2
        g = torch.Generator().manual_seed(0)
        loc = torch.nn.functional.normalize(torch.randn(n, d, generator=g), dim=-1)
        kappa = torch.exp(torch.rand(n, generator=g) * 5.3 + 2.3)
        ls = torch.tensor(LOGIT_SCALE)

        def step():
            li, lt = loc.clone().requires_grad_(True), loc.clone().requires_grad_(True)
            k = kappa.clone().requires_grad_(True)
            a, b = DO.ps_alpha_beta(d, k)
            z = torch.distributions.Beta(a, b).rsample((Ss,))
            v = torch.randn(Ss, n, d - 1)
            si, st = DO.ps_rsample(li, k, z, v), DO.ps_rsample(lt, k, z, v)
            kl = DO.ps_kl_uniform(d, k).mean()
            DO.vclip_total(si, st, kl, kl, 1.0 / k, 1.0 / k, ls, label_smoothing=0.0)[
                "total_loss"].backward()
        return (step, float(n) * Ss / S,
                "S=%d of %d sample slices (PowerSpherical restatement + clip_loss_sampled + KL), "
                "B=%d fp32: %.2f of one step" % (Ss, S, n, Ss / S))
    # alignclip: the N x N work cannot run whole on the host (N=65536: 17 GB per logits matrix);
    # a reduced batch is timed and scaled by the N^2 / N'^2 work ratio
    ns = min(n, 4096)
    x, y = synth_pair(ns, d, 1234)
    sem = torch.randn(ns, c["dsem"], generator=torch.Generator().manual_seed(5))
    sc = torch.tensor(14.285)

    def step():
        O.loss_and_grads(lambda a, b: O.inmodality_loss(a, b, sc, sem, 1.0, 0.5, True, True, False)[0],
                         [x, y])
    return (step, float(ns) * ns / n,
            "reduced batch N'=%d (oracle.inmodality_loss fwd+bwd, semantic branch, fp32), scaled "
            "by the work ratio (N'/N)^2 = %.5f of one step" % (ns, (ns / n) ** 2))


def cpu_baseline(args, reps=3):
    import torch
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    step, pairs, sample = cpu_case(args)
    step()  # warm-up (thread pool, allocator)
    t = min(_best(step, reps))
    out = {"value": pairs / t, "unit": "pairs/s", "cores": threads, "kind": "port",
           "sample": sample + ", best of %d" % reps, "seconds_per_sample": t}
    if args.cfg["loss"] == "clip" and args.cfg["n"] > 8192:
        # anchors for the slab extrapolation: cfg1 whole, and N=8192 whole vs its own slab
        f = cpu_clip_full(256, args.cfg["d"]); f()
        out["cfg1_n256_whole_pairs_per_s"] = 256 / min(_best(f, 5))
        f = cpu_clip_full(8192, args.cfg["d"]); f()
        tw = min(_best(f, 2))
        s = cpu_clip_slab(8192, args.cfg["d"], 1024); s()
        ts = min(_best(s, 2))
        out["n8192_whole_pairs_per_s"] = 8192 / tw
        out["n8192_slab1024_pairs_per_s"] = 1024 / ts
        out["slab_vs_whole_at_n8192"] = (1024 / ts) / (8192 / tw)
    return out


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (pinned oracle port) on the host cores."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    step, pairs, sample = cpu_case(args)
    for _ in range(max(1, args.warmup)):
        step()
    ts = _best(step, args.steps)
    t = sum(ts) / len(ts)
    value = pairs / t
    out = {
        "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3 * args.cfg["n"] / pairs, "ms_per_sample": t * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": workload_name(args.cfg, args), "config": args.config,
                   "global_batch": args.cfg["n"], "dim": args.cfg["d"]},
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def build_case(args, dev, rank, world):
    """The loss module, its pinned host inputs (this rank's shard) and the step function."""
    import torch
    import torch.nn.functional as F
    c = args.cfg
    n, d = c["n"], c["d"]
    n_local = n // world if c["scaling"] == "strong" else n
    sl = slice(rank * n_local, (rank + 1) * n_local) if c["scaling"] == "strong" else slice(0, n)
    x32, y32 = synth_pair(n, d, 1234)
    in_dtype = torch.bfloat16 if c["dtype"] == "bf16" else torch.float32
    case = {"n_local": n_local, "graph": not args.no_graph, "consts": []}
    if c["loss"] == "clip":
        hosts = [x32[sl].to(in_dtype), y32[sl].to(in_dtype)]
        ls = torch.tensor(LOGIT_SCALE, device=dev, requires_grad=True)
        if world > 1:
            from multimodal_2025_b200.dist import ShardedClipLoss
            crit = ShardedClipLoss(0.1).to(dev)
        else:
            from multimodal_2025_b200.losses import ClipLoss
            crit = ClipLoss(0.1).to(dev)
        case.update(crit=crit, ls=ls, fn=lambda a, b: crit(a, b, ls), clip_like=True)
    elif c["loss"] == "cloob":
        from multimodal_2025_b200.losses import CLOOBLoss
        from multimodal_2025_b200.losses import cloobLoss as cl
        if args.precision == "bf16":
            cl.COMPUTE_DTYPE = torch.bfloat16
        crit = CLOOBLoss(30.0, 15.0, dev)
        hosts = [x32.clone(), y32.clone()]
        case.update(crit=crit, fn=lambda a, b: crit(a, b))
    elif c["loss"] == "vclip":
        from multimodal_2025_b200.losses.distributions import (PowerSpherical, ProjectedNormal,
                                                               VonMisesFisher)
        from multimodal_2025_b200.losses.vclipLoss import VClipLoss
        S = c["S"]
        crit = VClipLoss(num_samples=S, label_smoothing=0.0).to(dev)
        if args.precision == "bf16":
            crit.clip_loss.compute_dtype = torch.bfloat16
        g = torch.Generator().manual_seed(0)
        kappa = torch.exp(torch.rand(n, generator=g) * 5.3 + 2.3).to(dev)   # LogUniform[10, 2000]
        sigma = torch.exp(torch.rand(n, d, generator=g) * 3.4 - 4.6).to(dev)
        ls = torch.tensor(LOGIT_SCALE, device=dev, requires_grad=True)
        hosts = [x32.clone(), y32.clone()]
        name = args.vdist

        def fn(a, b):
            k = kappa.clone().requires_grad_(True)
            if name == "power_spherical":
                di, dt_ = PowerSpherical(a, k), PowerSpherical(b, k)
            elif name == "vmf":
                di, dt_ = VonMisesFisher(a, k), VonMisesFisher(b, k)
            else:
                di, dt_ = ProjectedNormal(a, sigma), ProjectedNormal(b, sigma)
            v = 1.0 / k if name != "projected_normal" else sigma.mean(-1)
            return crit(di, dt_, v, v, ls, True)["total_loss"]
        # fresh Philox noise per call: the seed is a host value, so the step is not graph-captured
        case.update(crit=crit, ls=ls, fn=fn, graph=False)
    else:
        sem = torch.randn(n, c["dsem"], generator=torch.Generator().manual_seed(5))[sl].to(dev)
        sc = torch.tensor(14.285, device=dev)
        hosts = [x32[sl].to(in_dtype), y32[sl].to(in_dtype)]
        if world > 1:
            from multimodal_2025_b200.dist import ShardedClipInModalityLoss
            crit = ShardedClipInModalityLoss(1.0, 0.5, True, True, False)
        else:
            from multimodal_2025_b200.losses import AlignCLIPLoss
            crit = AlignCLIPLoss()  # plain class with .forward only (reference quirk D.2)
        case.update(crit=crit, fn=lambda a, b: crit.forward(a, b, sc, semantic_features=sem))
    case["hosts"] = [h.contiguous().pin_memory() for h in hosts]
    return case


def run_ours(args):
    import torch
    import torch.distributed as dist
    from multimodal_2025_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (our arm) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()
    assert L.mm25_device_ok() == 1, "not an sm_100 device"

    c = args.cfg
    case = build_case(args, dev, rank, world)
    hosts = case["hosts"]
    ins = [h.to(dev).requires_grad_(True) for h in hosts]
    fn = case["fn"]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def eager_step(tensors):
        for t in tensors:
            t.grad = None
        if case.get("ls") is not None:
            case["ls"].grad = None
        loss = fn(*tensors)
        loss.backward()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(f, k):
        """k steps, each bracketed by CUDA events after an L2 flush; returns summed ms."""
        evs = []
        for _ in range(k):
            flush.zero_()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            f()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    for _ in range(max(args.warmup, 3)):
        eager_step(ins)
    barrier()
    # one CUDA graph of the whole fwd+bwd (multimodal_2025_b200.graphs); --no-graph = eager
    graphed, graph_note, launches_per_step = None, "eager launches", None
    if case["graph"]:
        try:
            from multimodal_2025_b200.graphs import GraphedLossStep, GraphedStep
            c0 = L.mm25_launch_count()
            if case.get("clip_like"):
                graphed = GraphedLossStep(case["crit"], ins[0], ins[1], case["ls"], warmup=1)
            else:
                graphed = GraphedStep(fn, ins, warmup=1)
            # launches recorded into the graph = launches of ONE eager step (warm-up 1 + capture)
            launches_per_step = (L.mm25_launch_count() - c0) // 2
            graph_note = "cuda-graph replay"
        except Exception as e:  # capture unsupported -> measure the eager path, say so
            graphed, graph_note = None, "eager (graph capture failed: %s)" % type(e).__name__
    elif c["loss"] == "vclip":
        graph_note = "eager launches (fresh Philox seed per step: not graph-captured)"
    run_step = (lambda: graphed()) if graphed is not None else (lambda: eager_step(ins))
    for _ in range(max(args.warmup, 3)):
        run_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    c0 = L.mm25_launch_count()
    ms = timed(run_step, args.steps)
    barrier()
    launches = launches_per_step if graphed is not None else \
        (L.mm25_launch_count() - c0) // args.steps
    clocks = sampler.stop() if rank == 0 else None
    loss_val = float(graphed.loss) if graphed is not None else float(eager_step(ins).detach())

    # ---- e2e: public API, host inputs, H2D + D2H inside the timed region -------------------
    dbuf = [torch.empty_like(h, device=dev) for h in hosts]
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    def e2e_step():
        if graphed is not None:  # H2D straight into the graph's static inputs, replay, D2H
            out = graphed(*hosts)
            loss_host.copy_(out[0], non_blocking=True)
            return
        for dst, src in zip(dbuf, hosts):
            dst.copy_(src, non_blocking=True)
        loss = eager_step([t.detach().requires_grad_(True) for t in dbuf])
        loss_host.copy_(loss.detach().float(), non_blocking=True)

    def e2e_pipelined(k):
        """k steps through GraphedLossStep's double-buffered input pipeline: the H2D of step
        t+1 runs on a copy stream under step t.  ONE event pair brackets all k steps (every
        H2D, L2 flush, replay and D2H is inside it); returns total ms."""
        cur = torch.cuda.current_stream()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        flush.zero_()
        e0.record(cur)
        graphed.prefetch(hosts[0], hosts[1], after=e0)
        for t in range(k):
            if t + 1 < k:
                graphed.prefetch(hosts[0], hosts[1])
            if t > 0:
                flush.zero_()
            loss, _, _, _ = graphed.run_prefetched()
            loss_host.copy_(loss, non_blocking=True)
        e1.record(cur)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    e2e_note = "serial: H2D -> step -> D2H per step"
    for _ in range(2):
        e2e_step()
    barrier()
    if graphed is not None and case.get("clip_like") and not args.serial_e2e:
        e2e_pipelined(2)
        barrier()
        ms_e2e = e2e_pipelined(args.steps)
        e2e_note = ("double-buffered: H2D of step t+1 on a copy stream under step t "
                    "(graphs.GraphedLossStep.prefetch); all copies inside the timed region")
    else:
        ms_e2e = timed(e2e_step, args.steps)
    barrier()

    # ---- per-kernel times INSIDE eager steps (library hook: CUDA events around each launch of
    #      the hot kernels, on the launch stream) ---------------------------------------------
    slots = None
    try:
        k = max(2, min(args.steps, 5))
        eager_step(ins)
        torch.cuda.synchronize()
        L.mm25_debug_timing(1)
        t_eager = timed(lambda: eager_step(ins), k) / k
        import ctypes
        slots = {}
        for s in range(8):
            tot, cnt = ctypes.c_double(0.0), ctypes.c_longlong(0)
            L.mm25_debug_timing_read(s, ctypes.byref(tot), ctypes.byref(cnt))
            if cnt.value:
                slots[s] = {"ms_per_step": tot.value / k, "launches_per_step": cnt.value / k}
        L.mm25_debug_timing(0)
        slots["eager_ms_per_step"] = t_eager
    except Exception as e:  # noqa: BLE001 - the step timing above must still be reported
        slots = {"error": repr(e)[:200]}
        L.mm25_debug_timing(0)
    barrier()

    coll_note = "none (single GPU)" if world == 1 else "replicas only (no collective)"
    if world > 1 and c["scaling"] == "strong":
        import multimodal_2025_b200.dist as mdist
        peer = any(k[0] == "yall" and v is not None for k, v in mdist._SYMM_CACHE.items()
                   if isinstance(k[0], str))
        coll_note = (("one-shot all-gather over NVLink peer memory (Yhat, col stats, scalar "
                      "partials) + symmetric-memory barriers, no NCCL in the step" if peer
                      else "nccl all-gather + all-gather(col stats) + all-reduce(scalars)")
                     + "; gathered-operand gradient: " + str(mdist.LAST_BACKWARD_PATH))
    tm = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(tm[0]), float(tm[1])
    ms_step = ms / args.steps
    replicas = world if c["scaling"] == "replicas" else 1
    pairs = c["n"] * replicas
    return dict(world=world, rank=rank, n=c["n"], d=c["d"], n_local=case["n_local"],
                ms_step=ms_step, value=pairs / (ms_step * 1e-3),
                e2e_value=pairs / (ms_e2e / args.steps * 1e-3), launches=int(launches),
                clocks=clocks, loss=loss_val, h2d=int(sum(h.numel() * h.element_size() for h in hosts)),
                dist=dist, graph_note=graph_note, coll_note=coll_note, e2e_note=e2e_note,
                slots=slots, replicas=replicas)


def slot_algorithmic(c, r):
    """Algorithmic FLOPs (or bytes for the samplers) per STEP and rank attributed to each timed
    kernel slot: the forward sweep owns the 2 M N d of the logits, the backward
    GEMMs (gemm_te or gemm_grouped) the 4 M N d of dA / dB; the S-tile recompute of
    sim_grad_tc_kernel is implementation overhead (0 algorithmic, executed 2 M N d)."""
    M, N, d = float(r["n_local"]), float(c["n"]), float(c["d"])
    if c["loss"] == "clip":
        return {0: 2 * M * N * d, 1: 4 * M * N * d, 2: 0.0, 3: 4 * M * N * d,
                5: 2 * M * N * d, 6: 4 * M * N * d}
    if c["loss"] == "vclip":
        S = c["S"]
        byt = 2 * (S * N * d * 4 + N * d * 4) + 2 * (2 * S * N * d * 4 + 2 * N * d * 4)
        return {0: S * 2 * N * N * d, 1: S * 4 * N * N * d, 2: 0.0, 3: S * 4 * N * N * d,
                5: S * 2 * N * N * d, 6: S * 4 * N * N * d, 7: float(byt)}
    if c["loss"] == "cloob":
        # 4 retrievals (QK^T 2, PK 2 | backward 8) + 2 InfoLOOB terms (2 | 4): SURVEY §8d 60 N^2 d
        return {0: 12 * N * N * d, 2: 0.0, 3: 24 * N * N * d, 4: 16 * N * N * d,
                5: 20 * N * N * d, 6: 40 * N * N * d}
    ds = float(c["dsem"])
    return {0: 2 * M * N * d, 1: 4 * M * N * d, 2: 0.0, 3: 4 * M * N * d,
            4: 2 * M * N * (d + ds), 5: 2 * M * N * d, 6: 4 * M * N * d}


def roofline_object(args, r, peaks):
    """The dominant kernel of the step from the in-step per-launch timings."""
    slots = r["slots"]
    if not slots or "error" in slots:
        return {"bound": "tensor", "error": (slots or {}).get("error", "no timing")}
    c = args.cfg
    algo = slot_algorithmic(c, r)
    kernels = {}
    for s, v in slots.items():
        if not isinstance(s, int):
            continue
        a = algo.get(s, 0.0)
        ent = {"ms_per_step": v["ms_per_step"], "launches_per_step": v["launches_per_step"],
               "ms_per_launch": v["ms_per_step"] / v["launches_per_step"]}
        if s == 7:
            gbs = a / (v["ms_per_step"] * 1e-3) / 1e9
            ent.update(bound="hbm", achieved_gbs=gbs, frac=gbs / peaks["hbm_gbs"],
                       algorithmic_bytes_per_step=a)
        else:
            tf = a / (v["ms_per_step"] * 1e-3) / 1e12
            ent.update(bound="tensor", algorithmic_tflops=tf, frac=tf / peaks["bf16_tflops"],
                       algorithmic_flops_per_step=a)
            if s == 2:
                ent["note"] = "S-tile recompute: executed FLOPs only, no algorithmic credit"
        kernels[SLOT_NAMES[s]] = ent
    dom = max((s for s in slots if isinstance(s, int)), key=lambda s: slots[s]["ms_per_step"])
    v, a = slots[dom], algo.get(dom, 0.0)
    name = SLOT_NAMES[dom]
    traffic, tsrc = None, None
    tpath = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        ent = tj.get("kernels", {}).get(name)
        if ent and tj.get("config") == args.config and tj.get("n_gpus") == r["world"] and \
                not args.n and not args.d:
            traffic = ent["dram_bytes_per_launch"]
            tsrc = tj.get("source")
    per_launch = a / v["launches_per_step"]
    ms_launch = v["ms_per_step"] / v["launches_per_step"]
    if dom == 7:
        ach = per_launch / (ms_launch * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": name, "achieved": ach, "peak": peaks["hbm_gbs"],
                "unit": "GB/s", "frac": ach / peaks["hbm_gbs"], "traffic": traffic,
                "traffic_source": tsrc, "peak_source": peaks["source"] + " hbm_gbs",
                "ms_per_launch": ms_launch, "algorithmic_bytes_per_launch": per_launch,
                "kernels": kernels, "eager_ms_per_step": slots.get("eager_ms_per_step")}
    ach = per_launch / (ms_launch * 1e-3) / 1e12
    return {"bound": "tensor", "kernel": name, "achieved": ach, "peak": peaks["bf16_tflops"],
            "unit": "TFLOP/s", "frac": ach / peaks["bf16_tflops"], "traffic": traffic,
            "traffic_source": tsrc,
            "peak_source": peaks["source"] + " bf16_tflops (burst: the timed region is < 1 s)",
            "ms_per_launch": ms_launch, "launches_per_step": v["launches_per_step"],
            "algorithmic_flops_per_launch": per_launch,
            "share_of_step": v["ms_per_step"] / slots["eager_ms_per_step"],
            "timing": "CUDA events around each launch inside eager steps (mm25_debug_timing), "
                      "L2 flushed before every step",
            "kernels": kernels, "eager_ms_per_step": slots.get("eager_ms_per_step")}


def main():
    args = parse_args()
    if args.impl == "reference":
        # under torchrun only rank 0 measures; the other ranks leave before importing torch so
        # that their start-up does not compete with the timed CPU run for the host cores
        if int(os.environ.get("RANK", "0")) != 0:
            return
        run_reference(args)
        return
    peaks = load_peaks()
    r = run_ours(args)
    roof = roofline_object(args, r, peaks)
    world, rank = r["world"], r["rank"]
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_baseline(args)
        except Exception as e:  # noqa: BLE001
            cpu = {"error": repr(e)[:200]}
    if world > 1:
        r["dist"].barrier()
    if rank == 0:
        c = args.cfg
        flops = algorithmic_flops(c) * r["replicas"]
        tf = flops / (r["ms_step"] * 1e-3) / 1e12 / world
        reduced = c["dtype"] == "f32" and args.precision == "bf16"
        out = {
            "metric": METRIC, "value": r["value"], "unit": "pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_step"],
            "higher_is_better": True,
            "scaling": "strong" if c["scaling"] == "strong" else "weak", "vs_baseline": None,
            "dtype": "bf16" if (c["dtype"] == "bf16" or reduced) else "f32", "data": "synthetic",
            "config": {"workload": workload_name(c, args), "config": args.config,
                       "global_batch": c["n"] * r["replicas"], "dim": c["d"],
                       "per_gpu_rows": r["n_local"],
                       "parallelism": ("single GPU" if world == 1 else
                                       "row-sharded x%d" % world if c["scaling"] == "strong"
                                       else "replicas only x%d" % world),
                       "l2_flush_between_steps": True, "launch": r["graph_note"],
                       "collectives": r["coll_note"]},
            "step_tflops_per_gpu_algorithmic": tf,
            "step_frac_of_bf16_peak": tf / peaks["bf16_tflops"],
            "step_peak": "%s bf16_tflops (burst) = %.1f" % (peaks["source"], peaks["bf16_tflops"]),
            "loss": r["loss"],
            "clocks": r["clocks"],
            "e2e": {"value": r["e2e_value"], "unit": "pairs/s",
                    "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": 4,
                    "pipeline": r["e2e_note"]},
            "gpu_launches": r["launches"],
            "roofline": roof,
        }
        if cpu is not None:
            out["cpu_baseline"] = cpu
        print(json.dumps(out))
    if world > 1:
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()

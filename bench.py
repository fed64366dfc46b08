#!/usr/bin/env python
"""bench.py — CLIP-loss fwd+bwd pairs/s at B=32k d=512 bf16 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 32768]

A "step" is one forward+backward of the contrastive loss over one synthetic batch of
L2-normalised bf16 embeddings.  `value` is timed with CUDA events per step with
inputs resident in HBM and an L2 flush before every step; `e2e` goes through the public
ClipLoss module with HOST (pinned) inputs, H2D inside the timed region and a D2H read of the loss.
`--impl reference` times the CPU restatement of the reference (oracle/) on the host cores, on a
bounded slab of the same workload.  Prints ONE JSON line on rank 0.
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


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--n", type=int, default=32768, help="global batch (pairs)")
    p.add_argument("--d", type=int, default=512)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-graph", action="store_true", help="eager launches instead of a CUDA graph")
    p.add_argument("--serial-e2e", action="store_true",
                   help="e2e without the double-buffered input pipeline (H2D, step, D2H in series)")
    p.add_argument("--slab-rows", type=int, default=1024, help="rows of the CPU sample slab")
    return p.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            pk = json.load(f)
        return {"bf16_tflops": pk.get("bf16_tflops", 1590.0),
                "bf16_tflops_sustained": pk.get("bf16_tflops_sustained", 1400.0),
                "hbm_gbs": pk.get("hbm_gbs", 6650.0), "source": "measured"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback"}


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
# CPU baseline: the oracle restatement (kind "port") on a bounded slab of the same workload
# ------------------------------------------------------------------------------------------
def cpu_slab_step(x, y, ls, rows, eps=0.1):
    """fwd+bwd of the CLIP loss terms owned by `rows` image rows against ALL N texts:
    logits slab [rows, N] (1/(N/rows) of the full N x N work, same per-pair work), row CE with
    label smoothing for those rows + the slab's share of the column CE, then backward.
    Follows oracle.losses_oracle.clip_loss (reference clipLoss.py:94-119) on the slab."""
    import torch
    import torch.nn.functional as F
    xs = x[:rows].detach().clone().requires_grad_(True)
    yy = y.detach().clone().requires_grad_(True)
    lsp = ls.detach().clone().requires_grad_(True)
    xn = F.normalize(xs, dim=-1, p=2)
    yn = F.normalize(yy, dim=-1, p=2)
    s = lsp.clamp(max=4.6052).exp()
    z = s.float() * xn.float() @ yn.t().float()                    # [rows, N]
    labels = torch.arange(rows)
    loss_img = F.cross_entropy(z, labels, label_smoothing=eps)
    col = torch.logsumexp(z, dim=0).mean() - z.diagonal().mean()   # slab share of the text CE
    loss = 0.5 * (loss_img + col)
    loss.backward()
    return float(loss.detach())


def time_cpu_slab(n, d, rows, reps, threads):
    import torch
    from oracle import losses_oracle as O
    torch.set_num_threads(threads)
    x, y = O.synth_embeddings(n, d, 1234)
    ls = torch.tensor(LOGIT_SCALE)
    cpu_slab_step(x, y, ls, min(rows, 256))  # warm-up (thread pool, allocator)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        cpu_slab_step(x, y, ls, rows)
        ts.append(time.perf_counter() - t0)
    return ts


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    rows = min(args.slab_rows, args.n)
    ts = time_cpu_slab(args.n, args.d, rows, args.warmup + args.steps, threads)[args.warmup:]
    t = sum(ts) / len(ts)
    frac = rows / float(args.n)
    value = rows / t   # pairs/s: a slab of `rows` pairs' rows costs rows/N of a full step
    out = {
        "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3 / frac,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": "CLIP InfoNCE fwd+bwd, global B=%d d=%d, label_smoothing 0.1"
                               % (args.n, args.d), "global_batch": args.n, "dim": args.d},
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": threads, "kind": "port",
                         "sample": "slab of %d of %d image rows x all %d texts (%.4f of one "
                                   "step), fp32, torch CPU" % (rows, args.n, args.n, frac)},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from oracle import losses_oracle as O
    import multimodal_2025_b200 as pkg
    from multimodal_2025_b200 import ops, _lib
    from multimodal_2025_b200.losses import ClipLoss

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

    n, d = args.n, args.d
    n_local = n // world
    # synthetic inputs: seeded on CPU, normalised, cast to bf16 AFTER normalising
    x32, y32 = O.synth_embeddings(n, d, 1234)
    sl = slice(rank * n_local, (rank + 1) * n_local)
    x_host = x32[sl].to(torch.bfloat16).contiguous().pin_memory()
    y_host = y32[sl].to(torch.bfloat16).contiguous().pin_memory()
    del x32, y32
    x = x_host.to(dev).requires_grad_(True)
    y = y_host.to(dev).requires_grad_(True)
    ls = torch.tensor(LOGIT_SCALE, device=dev, requires_grad=True)

    if world > 1:
        from multimodal_2025_b200.dist import ShardedClipLoss
        crit = ShardedClipLoss(0.1).to(dev)
    else:
        crit = ClipLoss(0.1).to(dev)

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step(xi, yi):
        xi.grad = None
        yi.grad = None
        ls.grad = None
        loss = crit(xi, yi, ls)
        loss.backward()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        """k steps, each bracketed by CUDA events after an L2 flush; returns summed ms."""
        evs = []
        for _ in range(k):
            flush.zero_()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    for _ in range(args.warmup):
        step(x, y)
    barrier()
    # one CUDA graph of the whole fwd+bwd (multimodal_2025_b200.graphs); --no-graph = eager
    graphed, graph_note = None, "eager"
    launches_per_step = None
    if not args.no_graph:
        try:
            from multimodal_2025_b200.graphs import GraphedLossStep
            c0 = L.mm25_launch_count()
            graphed = GraphedLossStep(crit, x, y, ls, warmup=1)
            # launches recorded into the graph = launches of ONE eager step (warm-up 1 + capture)
            launches_per_step = (L.mm25_launch_count() - c0) // 2
            graph_note = "cuda-graph replay"
        except Exception as e:  # capture unsupported -> measure the eager path, say so
            graphed, graph_note = None, "eager (graph capture failed: %s)" % type(e).__name__
    run_step = (lambda: graphed()) if graphed is not None else (lambda: step(x, y))
    for _ in range(args.warmup):
        run_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    c0 = L.mm25_launch_count()
    t_wall0 = time.perf_counter()
    ms = timed(run_step, args.steps)
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    launches = launches_per_step if graphed is not None else \
        (L.mm25_launch_count() - c0) // args.steps
    clocks = sampler.stop() if rank == 0 else None
    loss_val = float(graphed.loss) if graphed is not None else float(crit(x, y, ls).detach())

    # ---- e2e: public API, host inputs, H2D + D2H inside the timed region -------------------
    xd = torch.empty_like(x_host, device=dev)
    yd = torch.empty_like(y_host, device=dev)
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    def e2e_step():
        if graphed is not None:  # H2D straight into the graph's static inputs, replay, D2H
            loss, _, _, _ = graphed(x_host, y_host)
            loss_host.copy_(loss, non_blocking=True)
            return
        xd.copy_(x_host, non_blocking=True)
        yd.copy_(y_host, non_blocking=True)
        xi = xd.detach().requires_grad_(True)
        yi = yd.detach().requires_grad_(True)
        loss = step(xi, yi)
        loss_host.copy_(loss.detach(), non_blocking=True)

    def e2e_pipelined(k):
        """k steps through GraphedLossStep's double-buffered input pipeline: the H2D of step
        t+1 runs on a copy stream under step t.  ONE event pair brackets all k steps (every
        H2D, L2 flush, replay and D2H is inside it); returns total ms."""
        cur = torch.cuda.current_stream()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        flush.zero_()
        e0.record(cur)
        graphed.prefetch(x_host, y_host, after=e0)
        for t in range(k):
            if t + 1 < k:
                graphed.prefetch(x_host, y_host)
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
    if graphed is not None and not args.serial_e2e:
        e2e_pipelined(2)
        barrier()
        ms_e2e = e2e_pipelined(args.steps)
        e2e_note = ("double-buffered: H2D of step t+1 on a copy stream under step t "
                    "(graphs.GraphedLossStep.prefetch); all copies inside the timed region")
    else:
        ms_e2e = timed(e2e_step, args.steps)
    barrier()

    coll_note = "none (single GPU)"
    if world > 1:
        import multimodal_2025_b200.dist as mdist
        peer = any(k[0] == "yall" and v is not None for k, v in mdist._SYMM_CACHE.items()
                   if isinstance(k[0], str))
        coll_note = (("one-shot all-gather over NVLink peer memory (Yhat, col stats, scalar "
                      "partials) + symmetric-memory barriers, no NCCL in the step" if peer
                      else "nccl all-gather(Yhat) + all-gather(col stats) + all-reduce(scalars)")
                     + "; dY: " + str(mdist.LAST_BACKWARD_PATH))
    tm = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(tm[0]), float(tm[1])
    ms_step = ms / args.steps
    value = n / (ms_step * 1e-3)
    e2e_value = n / (ms_e2e / args.steps * 1e-3)
    return dict(world=world, rank=rank, dev=dev, n=n, d=d, n_local=n_local, x=x, y=y, ls=ls,
                ms_step=ms_step, value=value, e2e_value=e2e_value, launches=int(launches),
                clocks=clocks, loss=loss_val, wall_ms=wall_ms, flush=flush, ops=ops,
                h2d=int(x_host.numel() * 2 * 2), dist=dist, graph_note=graph_note,
                coll_note=coll_note, e2e_note=e2e_note)


def kernel_roofline(r, args, peaks):
    """Per-kernel timing of the two fused kernels (CUDA events on the launch stream, L2 flushed
    before each launch) -> roofline object of the dominant one."""
    import torch
    ops = r["ops"]
    x, y, ls, flush = r["x"], r["y"], r["ls"], r["flush"]
    M, N, d = r["n_local"], r["n"], r["d"]
    with torch.no_grad():
        xh, _, _ = ops.rownorm_fwd(x.detach().unsqueeze(0), True, 1e-12, ops.BF16, False)
        yh, _, _ = ops.rownorm_fwd(y.detach().unsqueeze(0), True, 1e-12, ops.BF16, False)
        if N != M:  # sharded run: rank-local rectangular block [M, N]
            yh = yh.repeat(1, N // M, 1).contiguous()
        sc = ls.detach().reshape(1).float()
        off = r["rank"] * M

        keep = 2 * M * N <= ops.keep_e_bytes_limit(M, N, d)  # as the loss modules decide
        kept = [None, None]

        def fwd():
            if keep:
                out = ops.sim_fwd_keep(xh, yh, sc, ops.SCALE_LOG, 4.6052, off, True,
                                       ops.IMPL_AUTO, ops.DOT_BOUND_NORMALIZED)
                kept[0], kept[1] = out[4], out[5]
                return out[:4]
            return ops.sim_fwd(xh, yh, None, None, None, sc, ops.SCALE_LOG, 4.6052,
                               ops.SIM_PLAIN, off, True, ops.IMPL_AUTO,
                               ops.DOT_BOUND_NORMALIZED)

        row_lse, cmax, csum, _ = fwd()
        col_lse = ops.col_merge(cmax.reshape(1, -1), csum.reshape(1, -1)).reshape(1, -1)
        nn_ = float(N)

        def bwd():
            return ops.sim_bwd(xh, yh, None, None, None, sc, ops.SCALE_LOG, 4.6052,
                               ops.SIM_PLAIN, off, row_lse, col_lse, 0.5 / nn_, 0.5 / nn_,
                               0.9 / nn_, 0.1 / (nn_ * nn_), None, True, True, False,
                               ops.IMPL_AUTO, ops.DOT_BOUND_NORMALIZED, kept[0], kept[1])

        def t(fn, k, prep=None):
            tot = 0.0
            for _ in range(k):
                if prep is not None:
                    prep()  # untimed: the backward consumes the E its forward kept
                flush.zero_()
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                tot += e0.elapsed_time(e1)
            return tot / k

        k = max(3, min(args.steps, 10))
        fwd(); bwd()
        ms_f, ms_b = t(fwd, k), t(bwd, k, fwd if keep else None)
        # The backward OP is several launches (per gradient panel: sim_grad_tc_kernel, then
        # gemm_grouped_kernel).  Time the grouped GEMM launch by itself, on the panel shapes the
        # op uses, through the library's test entry; a failure here only drops the extra keys.
        gemm = None
        try:
            gemm = time_grouped_gemm(r, xh, yh, flush, k)
        except Exception as e:  # noqa: BLE001 - never let the optional timing break the bench
            gemm = {"error": repr(e)[:200]}
    fl_f, fl_b = 2.0 * M * N * d, 4.0 * M * N * d
    tf_f, tf_b = fl_f / (ms_f * 1e-3) / 1e12, fl_b / (ms_b * 1e-3) / 1e12
    peak = peaks["bf16_tflops"]
    dom = ("sim_bwd", tf_b, ms_b, fl_b) if ms_b >= ms_f else ("sim_fwd", tf_f, ms_f, fl_f)
    # DRAM bytes per launch of the dominant op from the committed ncu --set full capture
    # (profiles/r1_v5_ncu_summary.md: dram read + write of each kernel; sim_bwd = 2 panels x
    # (grad 0.252 + 1.022 GB, grouped GEMM 2.342 + 0.087 GB), sim_fwd 0.067 + 0.059 GB).  Only for
    # the shape that capture was taken on.
    traffic = None
    if M == 32768 and N == 32768 and d == 512 and not keep:
        traffic = {"sim_bwd": 2 * (0.252 + 1.022 + 2.342 + 0.087) * 1e9,
                   "sim_fwd": (0.067 + 0.059) * 1e9}[dom[0]]
    traffic_src = "profiles/r1_v5_ncu_summary.md (ncu --set full, bytes per launch)"
    extra = {}
    # The dominant KERNEL (largest share of the step) is the grouped GEMM: `panels` launches per
    # step.  When its stand-alone timing succeeded it is the headline object; the fused-op
    # figures (which charge the backward op for its recompute) stay under "kernels".
    if gemm and "ms" in gemm and gemm["launches_per_step"] * gemm["ms"] >= max(
            ms_f, ms_b - gemm["launches_per_step"] * gemm["ms"]):
        tf_g = gemm["flops"] / (gemm["ms"] * 1e-3) / 1e12
        extra = {"op_level": {"kernel": dom[0], "achieved": dom[1], "frac": dom[1] / peak,
                              "ms": dom[2], "algorithmic_flops": dom[3],
                              "note": "whole fused op (all its launches) against its "
                                      "algorithmic FLOPs, recompute not credited"},
                 "share_of_step_ms": gemm["launches_per_step"] * gemm["ms"]}
        dom = ("gemm_grouped_kernel", tf_g, gemm["ms"], gemm["flops"])
        traffic = None
        if M == 32768 and N == 32768 and d == 512:
            traffic = (2.275 + 0.085) * 1e9
            traffic_src = "profiles/r1_final_ncu_summary.md (ncu --set full, bytes per launch)"
    return {
        "bound": "tensor", "kernel": dom[0], "achieved": dom[1], "peak": peak,
        "unit": "TFLOP/s", "frac": dom[1] / peak, "traffic": traffic,
        "traffic_source": traffic_src if traffic else None,
        "peak_source": peaks["source"] + " bf16_tflops (burst: kernel timed alone)",
        "ms_per_launch": dom[2], "algorithmic_flops_per_launch": dom[3], **extra,
        "kernels": {"sim_fwd": {"ms": ms_f, "tflops": tf_f, "frac": tf_f / peak,
                                "flops": fl_f},
                    "sim_bwd": {"ms": ms_b, "tflops": tf_b, "frac": tf_b / peak,
                                "flops": fl_b},
                    **per_launch_kernels(gemm, ms_b, peak)},
    }


def time_grouped_gemm(r, xh, yh, flush, k):
    """One `gemm_grouped_kernel` launch on the shapes of one gradient panel of the backward
    (dA[R, d] = G'[R, N] Yhat, dB[N, d] = G'^T Xhat[R]), timed alone with CUDA events on the
    launch stream after an L2 flush.  G' is synthetic (the GEMM's time does not depend on the
    values).  Uses the library's test entry mm25_debug_gemm_grouped (include/mm25.h)."""
    import torch
    from multimodal_2025_b200 import _lib
    M, N, d = r["n_local"], r["n"], r["d"]
    R = min(M, max(128, ((1 << 30) // (2 * N)) // 128 * 128))  # rows per 1 GiB bf16 panel
    panels = (M + R - 1) // R
    dev = xh.device
    G = (torch.randn(R, N, device=dev, dtype=torch.float32) * 0.01).to(torch.bfloat16)
    Y = yh[0].contiguous()
    X = xh[0, :R].contiguous()
    dA = torch.empty(R, d, device=dev, dtype=torch.float32)
    dB = torch.empty(N, d, device=dev, dtype=torch.float32)
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream

    def launch():
        rc = L.mm25_debug_gemm_grouped(G.data_ptr(), 0, Y.data_ptr(), 1, dA.data_ptr(), R, N,
                                       G.data_ptr(), 1, X.data_ptr(), 1, dB.data_ptr(), N, R, d,
                                       st)
        if rc != 0:
            raise RuntimeError("mm25_debug_gemm_grouped rc=%d" % rc)

    launch()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(k):
        flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        launch()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return {"ms": tot / k, "flops": 4.0 * R * N * d, "launches_per_step": panels,
            "panel_rows": R}


def per_launch_kernels(gemm, ms_bwd_op, peak):
    """Entries for the two kernels inside the backward op: the grouped GEMM as measured alone,
    the gradient-panel kernel as the remainder of the op (derived, not measured alone)."""
    if not gemm or "ms" not in gemm:
        return {"gemm_grouped_kernel": gemm} if gemm else {}
    tf = gemm["flops"] / (gemm["ms"] * 1e-3) / 1e12
    p = gemm["launches_per_step"]
    out = {"gemm_grouped_kernel": {
        "ms": gemm["ms"], "tflops": tf, "frac": tf / peak, "flops": gemm["flops"],
        "launches_per_step": p, "panel_rows": gemm["panel_rows"],
        "note": "one launch = dA and dB of one gradient panel; algorithmic = executed FLOPs; "
                "timed alone through mm25_debug_gemm_grouped on synthetic G'"}}
    rest = (ms_bwd_op - p * gemm["ms"]) / p
    if rest > 0:
        fl = gemm["flops"] / 2.0  # recompute of the panel's logits: 2 R N d (not algorithmic)
        out["sim_grad_tc_kernel"] = {
            "ms": rest, "executed_tflops": fl / (rest * 1e-3) / 1e12, "launches_per_step": p,
            "note": "derived: (backward op - grouped GEMM launches) / panels; recompute FLOPs "
                    "are implementation overhead, so no algorithmic rate is quoted"}
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    peaks = load_peaks()
    r = run_ours(args)
    try:
        roof = kernel_roofline(r, args, peaks)
    except Exception as e:  # noqa: BLE001 - the step timing above must still be reported
        roof = {"bound": "tensor", "error": repr(e)[:300]}
    world, rank = r["world"], r["rank"]
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        rows = min(args.slab_rows, args.n)
        ts = time_cpu_slab(args.n, args.d, rows, 3, threads)
        tbest = min(ts)
        cpu = {"value": rows / tbest, "unit": "pairs/s", "cores": threads, "kind": "port",
               "sample": "slab of %d of %d image rows x all %d texts (%.4f of one step), fp32 "
                         "torch CPU, best of 3" % (rows, args.n, args.n, rows / args.n)}
    if world > 1:
        r["dist"].barrier()
    if rank == 0:
        flops = 6.0 * r["n"] * r["n"] * r["d"]
        tf = flops / (r["ms_step"] * 1e-3) / 1e12 / world
        out = {
            "metric": METRIC, "value": r["value"], "unit": "pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "CLIP InfoNCE fwd+bwd, global B=%d d=%d bf16, "
                                   "label_smoothing 0.1, logit_scale log(1/0.07)"
                                   % (r["n"], r["d"]),
                       "global_batch": r["n"], "dim": r["d"], "per_gpu_rows": r["n_local"],
                       "parallelism": "row-sharded x%d" % world if world > 1 else "single GPU",
                       "l2_flush_between_steps": True, "launch": r["graph_note"],
                       "collectives": r["coll_note"]},
            "step_tflops_per_gpu_algorithmic_6N2d": tf,
            "step_frac_of_bf16_peak": tf / peaks["bf16_tflops_sustained"],
            "step_peak": "%s bf16_tflops_sustained" % peaks["source"],
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

#!/usr/bin/env python
"""MODULATED (AlignCLIP semantic term) kernels on one B200 at the cfg5 per-GPU block:
local rows M = N / P of the N = 65536 global batch, d = 1024, d_sem = 768, bf16.
Prints one JSON line per shape: forward / backward ms and TFLOP/s against the algorithmic count
(fwd 2 M N (d + d2); bwd 2 M N (d + d2) recompute is NOT counted, dA + dB = 4 M N d)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import multimodal_2025_b200  # noqa: E402,F401
from multimodal_2025_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def t(fn, k=5, warm=2):
    for _ in range(warm):
        fn()
    tot = 0.0
    for _ in range(k):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / k


def run(M, N, d, d2, impl):
    g = torch.Generator().manual_seed(0)
    nrm = torch.nn.functional.normalize
    b = nrm(torch.randn(1, N, d, generator=g), dim=-1).to(dev).bfloat16()
    a = b[:, :M].contiguous()
    v = nrm(torch.randn(1, N, d2, generator=g), dim=-1).to(dev)
    u = v[:, :M].contiguous()
    D = torch.rand(1, M, generator=g).to(dev)
    sc = torch.tensor([14.285], device=dev)
    fwd = lambda: ops.sim_fwd(a, b, u, v, D, sc, ops.SCALE_MUL, 0.0, ops.SIM_MODULATED, 0, False,
                              impl)
    row_lse = fwd()[0]
    path_f = _lib.lib().mm25_debug_last_sim_path()
    bwd = lambda: ops.sim_bwd(a, b, u, v, D, sc, ops.SCALE_MUL, 0.0, ops.SIM_MODULATED, 0,
                              row_lse, None, 1.0 / N, 0.0, 1.0 / N, 0.0, None, True, True, True,
                              impl)
    bwd()
    path_b = _lib.lib().mm25_debug_last_sim_path()
    k = 5 if impl != ops.IMPL_SIMT else 1
    ms_f, ms_b = t(fwd, k, 1), t(bwd, k, 1)
    fl_f, fl_b = 2.0 * M * N * (d + d2), 4.0 * M * N * d
    print(json.dumps({"path": "MODULATED sim_fwd/sim_bwd", "M": M, "N": N, "d": d, "d2": d2,
                      "impl": {0: "auto", 1: "simt"}[impl], "kernel_family": [path_f, path_b],
                      "fwd_ms": ms_f, "bwd_ms": ms_b, "fwd_tflops": fl_f / ms_f / 1e9,
                      "bwd_tflops_algorithmic": fl_b / ms_b / 1e9,
                      "step_tflops_algorithmic": (fl_f + fl_b) / (ms_f + ms_b) / 1e9}), flush=True)


if __name__ == "__main__":
    run(8192, 65536, 1024, 768, ops.IMPL_AUTO)      # cfg5 per-GPU block at P = 8
    if "--only-cfg5" not in sys.argv:
        run(8192, 8192, 1024, 768, ops.IMPL_AUTO)   # single-GPU square (profiles/r1_paths.jsonl)
        run(8192, 8192, 1024, 768, ops.IMPL_SIMT)   # the fp32 SIMT kernels on the same input

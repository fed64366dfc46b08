"""Timing of the kept-E backward GEMMs (csrc/gemm_te.cu) in isolation, through the C-ABI test
entry: both problems, dA only, dB only, copy mode (flag 0), and the TMA-fed grouped GEMM on the
same shapes for reference.  python tools/bench_te.py [N] [d]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from multimodal_2025_b200 import _lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
d = int(sys.argv[2]) if len(sys.argv) > 2 else 512
M = int(sys.argv[3]) if len(sys.argv) > 3 else N
L = _lib.lib()
dev = "cuda"
E = (torch.rand(M, N, device=dev) * 0.5).bfloat16()
rowf = torch.rand(M, device=dev) * 1e-3
colf = torch.rand(N, device=dev) * 1e-3
consts = torch.tensor([1e-6, 0.0], device=dev)
gdiag = torch.rand(M, device=dev) * 1e-4
one = torch.ones(1, dtype=torch.int32, device=dev)
zero = torch.zeros(1, dtype=torch.int32, device=dev)
Y = torch.randn(N, d, device=dev).bfloat16()
X = torch.randn(M, d, device=dev).bfloat16()
C0 = torch.empty(M, d, device=dev)
C1 = torch.empty(N, d, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream


def te(flag, a=True, b=True):
    rc = L.mm25_debug_gemm_te(E.data_ptr(), M, N, rowf.data_ptr(), colf.data_ptr(),
                              consts.data_ptr(), gdiag.data_ptr(), flag.data_ptr(), 0,
                              Y.data_ptr() if a else None, C0.data_ptr() if a else None,
                              X.data_ptr() if b else None, C1.data_ptr() if b else None, d, 1, st)
    assert rc == 0, rc


def grouped():
    rc = L.mm25_debug_gemm_grouped(E.data_ptr(), 0, Y.data_ptr(), 1, C0.data_ptr(), M, N,
                                   E.data_ptr(), 1, X.data_ptr(), 1, C1.data_ptr(), N, M, d, st)
    assert rc == 0, rc


def timeit(fn, k=5):
    fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(k):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / k


fl = 2.0 * M * N * d
import ctypes
STATS = os.environ.get('MM25_TE_STATS') == '1'
if STATS:
    L.mm25_debug_te_stats(None, 1)
for name, fn, f in (("te both (transform)", lambda: te(one), 2 * fl),
                    ("te dA only", lambda: te(one, True, False), fl),
                    ("te dB only", lambda: te(one, False, True), fl),
                    ("te both (copy mode)", lambda: te(zero), 2 * fl),
                    ("grouped TMA 1-CTA / 2-CTA per env", grouped, 2 * fl)):
    ms = timeit(fn)
    if STATS:
        buf = (ctypes.c_longlong * 8)()
        L.mm25_debug_te_stats(ctypes.cast(buf, ctypes.c_void_p), 1)
        n = max(buf[3], 1)
        print(json.dumps({"path": name, "cycles_per_tile": {"load_wait": buf[0] / n,
              "stage_wait": buf[1] / n, "fill": buf[2] / n, "tiles": buf[3]},
              "sm_ghz": buf[4] / max(buf[5], 1)}))
    print(json.dumps({"path": name, "M": M, "N": N, "d": d, "ms": ms,
                      "tflops": f / ms / 1e9}))

"""time the per-rank kernels of a P-way row shard on ONE GPU: local block [N/P, N]."""
import sys, os, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import multimodal_2025_b200
from multimodal_2025_b200 import ops
N, d = 32768, 512
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(fn, k=8):
    for _ in range(2): fn()
    tot = 0
    for _ in range(k):
        flush.zero_(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); tot += e0.elapsed_time(e1)
    return tot / k
for P in (1, 2, 4, 8):
    M = N // P
    g = torch.Generator().manual_seed(0)
    a = torch.nn.functional.normalize(torch.randn(1, M, d, generator=g), dim=-1).cuda().bfloat16()
    b = torch.nn.functional.normalize(torch.randn(1, N, d, generator=g), dim=-1).cuda().bfloat16()
    sc = torch.tensor([2.6592], device="cuda")
    fwd = lambda: ops.sim_fwd(a, b, None, None, None, sc, 1, 4.6052, 0, 0, True, 0, 1.01)
    r, cm, cs, _ = fwd()
    cl = ops.col_merge(cm.reshape(1, -1), cs.reshape(1, -1)).reshape(1, -1)
    nn = float(N)
    bwd = lambda: ops.sim_bwd(a, b, None, None, None, sc, 1, 4.6052, 0, 0, r, cl, .5/nn, .5/nn, .9/nn, .1/nn/nn, None, True, True, False, 0, 1.01)
    print("P=%d M=%d fwd %.3f ms bwd %.3f ms  (ideal 1/P of P=1)" % (P, M, t(fwd), t(bwd)))

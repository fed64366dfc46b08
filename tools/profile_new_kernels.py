"""Workload for `ncu -k regex:...` captures of the round-2 kernels outside the cfg2 step:
fused rank sweep, sliced top-k, fp32-grade split path (forward sweep K = 4 d, panel sweep,
grouped GEMM), Hopfield dual-accumulator sweep on split operands."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from multimodal_2025_b200 import ops  # noqa: E402
from multimodal_2025_b200.losses import CLOOBLoss, ClipLoss  # noqa: E402
from multimodal_2025_b200.metrics.retrieval import fused_ranks  # noqa: E402

g = torch.Generator().manual_seed(0)
dev = "cuda"
n, d = 32768, 512
x = torch.nn.functional.normalize(torch.randn(n, d, generator=g), dim=-1).to(dev)
y = torch.nn.functional.normalize(x + 0.05 * torch.randn(n, d, device=dev), dim=-1)
ar = torch.arange(n, device=dev)
for _ in range(2):
    fused_ranks(x.bfloat16(), y.bfloat16(), ar)
    ops.topk_rows(torch.randn(1, 1 << 20, generator=g).to(dev), 5)
    ops.topk_rows(torch.randn(300, 16384, generator=g).to(dev), 100)
    a = x[:8192].clone().requires_grad_(True)
    b = y[:8192].clone().requires_grad_(True)
    ls = torch.tensor(2.6592, device=dev, requires_grad=True)
    ClipLoss(0.1).to(dev)(a, b, ls).backward()
    a = x[:4096].clone().requires_grad_(True)
    b = y[:4096].clone().requires_grad_(True)
    CLOOBLoss(30.0, 15.0, dev)(a, b).backward()
torch.cuda.synchronize()
print("done")

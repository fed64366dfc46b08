"""CUDA-graph capture of one loss forward+backward.

The fused kernels are long (0.1-1 ms) but a step also launches ~25 small ones (normalise, merges,
reductions, collectives); from Python their launch gaps cost 5-10 % of a single-GPU step and far
more once the batch is sharded 8 ways.  `GraphedLossStep` captures `loss = crit(x, y, scale);
loss.backward()` once on static buffers (the kernels never allocate and take all sizes from the
host at capture time, so the captured launches stay valid) and replays it with one
`cudaGraphLaunch`.  NCCL collectives issued by `dist.ShardedClipLoss` are captured with it.

    step = GraphedLossStep(ClipLoss(0.1).cuda(), x, y, logit_scale)
    loss, dx, dy, dscale = step(x_new, y_new)      # copies into the static inputs, replays
"""
from __future__ import annotations

import torch


class GraphedLossStep:
    def __init__(self, crit, x, y, logit_scale, warmup: int = 3):
        if not x.is_cuda:
            raise RuntimeError("GraphedLossStep needs CUDA tensors (no CPU fallback)")
        self.crit = crit
        self.x = x.detach().clone().requires_grad_(True)
        self.y = y.detach().clone().requires_grad_(True)
        self.scale = logit_scale.detach().clone().requires_grad_(True)
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.x.grad = self.y.grad = self.scale.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            loss = crit(self.x, self.y, self.scale)
            loss.backward()
            self.loss = loss.detach()
        torch.cuda.synchronize()

    def _eager(self):
        self.x.grad = self.y.grad = self.scale.grad = None
        loss = self.crit(self.x, self.y, self.scale)
        loss.backward()
        return loss

    def __call__(self, x=None, y=None, logit_scale=None):
        """Replay on new inputs (copied into the static buffers; None = keep the current ones).
        Returns (loss, dL/dx, dL/dy, dL/dscale) — static tensors overwritten by the next call."""
        with torch.no_grad():
            if x is not None:
                self.x.copy_(x, non_blocking=True)
            if y is not None:
                self.y.copy_(y, non_blocking=True)
            if logit_scale is not None:
                self.scale.copy_(logit_scale, non_blocking=True)
        self.graph.replay()
        return self.loss, self.x.grad, self.y.grad, self.scale.grad

"""CUDA-graph capture of one loss forward+backward.

The fused kernels are long (0.1-1 ms) but a step also launches ~25 small ones (normalise, merges,
reductions, collectives); from Python their launch gaps cost 5-10 % of a single-GPU step and far
more once the batch is sharded 8 ways.  `GraphedLossStep` captures `loss = crit(x, y, scale);
loss.backward()` once on static buffers (the kernels never allocate and take all sizes from the
host at capture time, so the captured launches stay valid) and replays it with one
`cudaGraphLaunch`.  NCCL collectives issued by `dist.ShardedClipLoss` are captured with it.

    step = GraphedLossStep(ClipLoss(0.1).cuda(), x, y, logit_scale)
    loss, dx, dy, dscale = step(x_new, y_new)      # copies into the static inputs, replays

Host-resident batches (a pinned-memory loader) can be double-buffered so that the host->device
copy of batch t+1 runs on a copy stream while batch t computes:

    step.prefetch(x_host_0, y_host_0)
    for t in range(T):
        if t + 1 < T:
            step.prefetch(x_host[t + 1], y_host[t + 1])   # overlaps with the replay below
        loss, dx, dy, dscale = step.run_prefetched()
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

    # ---- double-buffered host -> device input pipeline ---------------------------------------
    def _ensure_pipeline(self):
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream()
            self._stage = [(torch.empty_like(self.x), torch.empty_like(self.y)) for _ in range(2)]
            self._ready = [torch.cuda.Event(), torch.cuda.Event()]
            self._free = [None, None]
            self._slot = 0
            self._queue = []

    def prefetch(self, x_host, y_host, after=None):
        """Enqueue the H2D copy of the next batch (pinned host tensors) on the copy stream into
        the idle staging slot; at most two batches may be in flight.  `after`: optional event the
        copy must not start before (benchmarks: the start-of-timing event)."""
        self._ensure_pipeline()
        if len(self._queue) >= 2:
            raise RuntimeError("prefetch: both staging slots are in flight; call run_prefetched()")
        s, cs = self._slot, self._copy_stream
        if after is not None:
            cs.wait_event(after)
        if self._free[s] is not None:      # the previous consumer of this slot has copied it out
            cs.wait_event(self._free[s])
        with torch.cuda.stream(cs), torch.no_grad():
            self._stage[s][0].copy_(x_host, non_blocking=True)
            self._stage[s][1].copy_(y_host, non_blocking=True)
            self._ready[s].record(cs)
        self._queue.append(s)
        self._slot ^= 1

    def run_prefetched(self):
        """Replay on the oldest prefetched batch; returns what __call__ returns."""
        if not getattr(self, "_queue", None):
            raise RuntimeError("run_prefetched: nothing was prefetched")
        s = self._queue.pop(0)
        cur = torch.cuda.current_stream()
        cur.wait_event(self._ready[s])
        with torch.no_grad():
            self.x.copy_(self._stage[s][0], non_blocking=True)
            self.y.copy_(self._stage[s][1], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(cur)
        self._free[s] = ev
        self.graph.replay()
        return self.loss, self.x.grad, self.y.grad, self.scale.grad

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


class GraphedStep:
    """CUDA-graph capture of `loss = fn(*inputs); loss.backward()` for any of the loss modules
    whose forward takes only tensors (CLOOBLoss, ClipInModalityLoss with semantic features, ...).
    Losses that draw fresh noise per call (VClipLoss: the Philox seed is a host value) must not be
    captured — the replay would reuse the captured seed.

        step = GraphedStep(lambda a, b: crit(a, b), [img, txt])
        loss, (dimg, dtxt) = step(img_new, txt_new)

    Inputs with `requires_grad=False` are treated as constants-with-storage (copied into, no grad).
    """

    def __init__(self, fn, inputs, warmup: int = 3):
        if not all(t.is_cuda for t in inputs):
            raise RuntimeError("GraphedStep needs CUDA tensors (no CPU fallback)")
        self.fn = fn
        self.inputs = [t.detach().clone().requires_grad_(t.requires_grad) for t in inputs]
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        for t in self.inputs:
            t.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            loss = fn(*self.inputs)
            loss.backward()
            self.loss = loss.detach()
        torch.cuda.synchronize()

    def _eager(self):
        for t in self.inputs:
            t.grad = None
        loss = self.fn(*self.inputs)
        loss.backward()
        return loss

    def __call__(self, *new_inputs):
        """Replay on new inputs (None / omitted = keep).  Returns (loss, [grad or None, ...]) —
        static tensors overwritten by the next call."""
        with torch.no_grad():
            for dst, src in zip(self.inputs, new_inputs):
                if src is not None:
                    dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.loss, [t.grad for t in self.inputs]

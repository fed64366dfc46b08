"""torchrun helper: per-kernel CUDA time of one sharded CLIP step (torch.profiler, rank 0)."""
import os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import synth_pair
from multimodal_2025_b200.dist import ShardedClipLoss

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
N, d = 32768, 512
n = N // world
x32, y32 = synth_pair(N, d, 1234)
sl = slice(rank * n, (rank + 1) * n)
x = x32[sl].bfloat16().to(dev).requires_grad_(True)
y = y32[sl].bfloat16().to(dev).requires_grad_(True)
ls = torch.tensor(2.6592, device=dev, requires_grad=True)
crit = ShardedClipLoss(0.1).to(dev)

def step():
    x.grad = y.grad = ls.grad = None
    loss = crit(x, y, ls); loss.backward(); return loss
for _ in range(5): step()
torch.cuda.synchronize(); dist.barrier()
if os.environ.get("MM25_PROFILE_GRAPH", "1") == "1":
    from multimodal_2025_b200.graphs import GraphedLossStep
    graphed = GraphedLossStep(crit, x, y, ls, warmup=1)
    step = lambda: graphed()
    for _ in range(5): step()
    torch.cuda.synchronize(); dist.barrier()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # last step only: split by the first rownorm kernel occurrences
    starts = [i for i, e in enumerate(evs) if "rownorm_f" in e.name]
    first = starts[-2] if len(starts) >= 2 else 0
    last = evs[first:]
    t0 = last[0].time_range.start
    busy = 0.0
    for e in last:
        print("%8.1f %8.1f  %s" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, e.name[:70]))
        busy += e.time_range.end - e.time_range.start
    print("span_us", last[-1].time_range.end - t0, "busy_us", busy)
sys.stdout.flush()
torch.cuda.synchronize()
os._exit(0)  # skip the process-group teardown: it hung after graph capture at 8 ranks

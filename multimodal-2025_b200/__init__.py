"""multimodal-2025_b200 — B200-native (sm_100a) drop-in for the contrastive-loss hot path of
neskech/Multimodal-2025 (`losses/`, `losses/distributions/`, `metrics/topk.py`).

Import as `multimodal_2025_b200` (see the loader module of that name at the repo root; the
directory name carries a hyphen).  Layout:
  csrc/      hand-written CUDA kernels + the C-ABI (include/mm25.h)
  _lib.py    nvcc build + ctypes loader (no fallback: raises if the library is missing)
  ops.py     torch.library custom ops `mm25::*`
  losses/    nn.Module mirrors of the reference losses (same signatures)
  metrics/   top-k retrieval (bit-exact indices)
  dist.py    row-sharded multi-GPU CLIP loss (NCCL all-gather / stat merge / reduce-scatter)
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
__version__ = "0.1.0"


def build(force: bool = False, verbose: bool = False) -> str:
    return _lib.build(force=force, verbose=verbose)

"""Mirrors of the reference `losses/` package (same class names and signatures)."""
from .clipLoss import ClipLoss  # noqa: F401
from .cloobLoss import CLOOBLoss, cloob_loss, infoloob_loss, hopfield_retrieval, hopfield  # noqa: F401
from .alignCLIPLoss import AlignCLIPLoss  # noqa: F401
from .align_clip_loss import ClipInModalityLoss  # noqa: F401
from .negatives_cache import cached_negatives_clip_loss  # noqa: F401

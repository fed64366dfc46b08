"""Mirrors of the reference `losses/` package (same class names and signatures)."""
from .clipLoss import ClipLoss  # noqa: F401
from .cloobLoss import CLOOBLoss, cloob_loss, infoloob_loss, hopfield_retrieval, hopfield  # noqa: F401

"""Mirrors of the reference `losses/` package (same class names and signatures)."""
from .clipLoss import ClipLoss  # noqa: F401

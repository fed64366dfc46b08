"""Drop-in for the reference `losses/alignCLIPLoss.py` (lines 1-8).

The reference class is a thin front for the in-modality loss of the AlignCLIP fork with the
paper's weights and the natural-language semantic supervision switched on.  It is a plain class
(neither an nn.Module nor callable — SURVEY quirk D.2), so only `.forward(...)` exists, and the
scale it receives is passed on untouched, i.e. it must ALREADY be exponentiated.
Everything numerical happens in `align_clip_loss.ClipInModalityLoss` (MODULATED kernels).
"""
from .align_clip_loss import ClipInModalityLoss

# cross-modal weight, in-modality weight (reference :4; the fork's defaults, loss.py:143-144)
_ALPHA, _BETA = 1.0, 0.5


class AlignCLIPLoss:
    """`AlignCLIPLoss().forward(image_features, text_features, logits_scale,
    semantic_features=None)` -> scalar loss (reference :2-8)."""

    def __init__(self):
        settings = dict(alpha=_ALPHA, beta=_BETA, nl_semantic_supervision=True)
        self.criterion = ClipInModalityLoss(**settings)

    def forward(self, image_features, text_features, logits_scale, semantic_features=None):
        loss_module = self.criterion
        return loss_module(image_features, text_features, logits_scale,
                           semantic_features=semantic_features)

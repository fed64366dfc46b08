"""Drop-in for the reference `losses/alignCLIPLoss.py` (lines 1-8): a plain class (not an
nn.Module, not callable — quirk D.2) that forwards to ClipInModalityLoss(alpha=1.0, beta=0.5,
nl_semantic_supervision=True).  `logits_scale` is passed through unchanged, i.e. ALREADY
exponentiated."""
from .align_clip_loss import ClipInModalityLoss


class AlignCLIPLoss:
    def __init__(self):
        self.criterion = ClipInModalityLoss(alpha=1.0, beta=0.5, nl_semantic_supervision=True)

    def forward(self, image_features, text_features, logits_scale, semantic_features=None):
        # Compute the alignment loss between image and text features
        return self.criterion(image_features, text_features, logits_scale,
                              semantic_features=semantic_features)

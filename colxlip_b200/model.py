"""Scoring method of the reference's ColXLIP model class on the B200 path.

The reference defines `ColXLIP.compute_colbert_similarity(self, token_image_features,
token_text_features)` (model.py:611-629): the same contraction + max over patches as the loss,
followed by a plain mean over all T text tokens.  The encoders themselves are out of scope
(SURVEY 2, rows 7-9); a ColXLIP model picks the method up through this mixin, or the function can be
called directly for retrieval scoring.
"""
from __future__ import annotations

from .loss import compute_colbert_similarity_mean


class ColXLIPScoringMixin:
    """Mix into (or monkey-patch onto) the reference's ColXLIP class."""

    def compute_colbert_similarity(self, token_image_features, token_text_features):
        """[Bt,Bi] plain-mean MaxSim scores, values in [-1, 1] for normalised features."""
        return compute_colbert_similarity_mean(token_image_features, token_text_features)


def patch_colxlip(model_cls):
    """Install the B200 scorer on the reference's model class: patch_colxlip(colxlip.model.ColXLIP)."""
    model_cls.compute_colbert_similarity = ColXLIPScoringMixin.compute_colbert_similarity
    return model_cls

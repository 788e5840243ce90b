"""colxlip_b200: B200-native (sm_100a) MaxSim contrastive scoring + loss, drop-in for
lezhang7/colxlip's `colxlip.loss` API.  CUDA only; there is no CPU fallback."""
from .loss import (ClipLoss, ColClipAccumulation, ColClipLoss, compute_colbert_similarity, compute_colbert_similarity_mean, gather_features,
                   get_precision, get_strict_autocast, set_precision, set_strict_autocast)
from .model import ColXLIPScoringMixin, TokenHead, encode_tokens, eot_valid_len, normalize_tokens, patch_colxlip
from .retrieval import compute_retrieval, maxsim_score_matrix, retrieval_on_split

__all__ = ["ClipLoss", "ColClipAccumulation", "ColClipLoss", "compute_colbert_similarity", "compute_colbert_similarity_mean", "gather_features",
           "ColXLIPScoringMixin", "patch_colxlip", "normalize_tokens", "TokenHead", "encode_tokens", "eot_valid_len", "compute_retrieval", "maxsim_score_matrix", "retrieval_on_split", "set_precision", "get_precision",
           "set_strict_autocast", "get_strict_autocast"]

"""bsc -- B200-native contrastive hot path of BIOSCAN-CLIP (loss + closest-embedding retrieval).

Host side in Python/PyTorch (device memory, streams, torch.distributed), arithmetic in libbsc_hot.so
(hand-written sm_100a CUDA behind the C ABI of include/bsc_hot.h).  No CPU fallback.
"""
from ._lib import BscError, LIB_PATH  # noqa: F401
from .extract import get_feature_and_label, get_features_and_label  # noqa: F401
from .loss import ClipLoss, ContrastiveLoss  # noqa: F401
from .retrieval import IndexFlatIP, make_prediction, merge_across_ranks  # noqa: F401
from .taxonomy import TaxonomyEval, top_k_macro_accuracy, top_k_micro_accuracy  # noqa: F401

__all__ = ["ContrastiveLoss", "ClipLoss", "IndexFlatIP", "make_prediction", "merge_across_ranks", "TaxonomyEval", "top_k_micro_accuracy",
           "top_k_macro_accuracy", "get_feature_and_label", "get_features_and_label", "BscError", "LIB_PATH"]

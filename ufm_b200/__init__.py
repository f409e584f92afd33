"""ufm_b200 — B200-native engine for UFM's dense-correspondence forward pass.

``UniFlowMatchConfidence`` / ``UniFlowMatchClassificationRefinement`` keep the reference's
``predict_correspondences_batched()`` API; the compute is hand-written sm_100a CUDA behind ``libufm_b200.so``.
"""
from .configs import CONFIGS, get_config  # noqa: F401
from .models import (  # noqa: F401
    UFMClassificationRefinementOutput,
    UFMFlowFieldOutput,
    UFMMaskFieldOutput,
    UFMOutputInterface,
    UniFlowMatchClassificationRefinement,
    UniFlowMatchConfidence,
)

"""B200-native TriangleMix prefill path (mask generator + masked GQA prefill attention).

Public names mirror the reference fork's modules:
  vllm/attention/trianglemix_attention.py -> .trianglemix_attention
  vllm/attention/ascend_trianglemix.py    -> .ascend_trianglemix
  vllm/attention/trianglemix_config.py    -> .trianglemix_config
  vLLM attention-backend plug-in slot     -> .backend
  Qwen3Attention.forward between the projections (models/qwen3.py:160-190) -> .qwen3_attention
"""
from .trianglemix_attention import (TriangleMixAnalyzer, TriangleMixAttention, TriangleMixConfig,
                                    TriangleMixMaskGenerator)
from .ascend_trianglemix import (AscendNPUTriangleMixOptimizer, AscendTriangleMixAttention,
                                 create_trianglemix_config_from_hf)
from .trianglemix_config import TriangleMixInferenceConfig
from .backend import (TriangleMixBackend, TriangleMixImpl, TriangleMixMetadata,
                      TriangleMixMetadataBuilder)
from .qwen3_attention import Qwen3TriangleMixAttention, compute_cos_sin_cache
from .layer_selection import TriangleLayerSelector, middle_qk_gradient
from ._lib import TmxError

__all__ = [
    "TriangleMixConfig", "TriangleMixMaskGenerator", "TriangleMixAttention", "TriangleMixAnalyzer",
    "AscendTriangleMixAttention", "AscendNPUTriangleMixOptimizer", "create_trianglemix_config_from_hf",
    "TriangleMixInferenceConfig", "TriangleMixBackend", "TriangleMixImpl", "TriangleMixMetadata",
    "TriangleMixMetadataBuilder", "Qwen3TriangleMixAttention", "compute_cos_sin_cache", "TmxError", "TriangleLayerSelector", "middle_qk_gradient",
]

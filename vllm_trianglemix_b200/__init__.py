"""B200-native TriangleMix prefill path (mask generator + masked GQA prefill attention).

Public names mirror the reference fork's modules:
  vllm/attention/trianglemix_attention.py -> .trianglemix_attention
  vllm/attention/ascend_trianglemix.py    -> .ascend_trianglemix
  vllm/attention/trianglemix_config.py    -> .trianglemix_config
  vLLM attention-backend plug-in slot     -> .backend
"""
from .trianglemix_attention import (TriangleMixAnalyzer, TriangleMixAttention, TriangleMixConfig,
                                    TriangleMixMaskGenerator)
from .ascend_trianglemix import (AscendNPUTriangleMixOptimizer, AscendTriangleMixAttention,
                                 create_trianglemix_config_from_hf)
from .trianglemix_config import TriangleMixInferenceConfig
from .backend import (TriangleMixBackend, TriangleMixImpl, TriangleMixMetadata,
                      TriangleMixMetadataBuilder)
from ._lib import TmxError

__all__ = [
    "TriangleMixConfig", "TriangleMixMaskGenerator", "TriangleMixAttention", "TriangleMixAnalyzer",
    "AscendTriangleMixAttention", "AscendNPUTriangleMixOptimizer", "create_trianglemix_config_from_hf",
    "TriangleMixInferenceConfig", "TriangleMixBackend", "TriangleMixImpl", "TriangleMixMetadata",
    "TriangleMixMetadataBuilder", "TmxError",
]

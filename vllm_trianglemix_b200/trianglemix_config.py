"""Drop-in for the reference's vllm/attention/trianglemix_config.py (user-facing presets)."""
from dataclasses import dataclass, field
from typing import List, Optional

from .trianglemix_attention import TriangleMixConfig


@dataclass
class TriangleMixInferenceConfig:
    """Mirror of reference trianglemix_config.py:11-81: same ten fields, same defaults."""

    enable_trianglemix: bool = False
    num_triangle_layers: Optional[int] = None
    triangle_layer_indices: List[int] = field(default_factory=list)
    num_sink_tokens: int = 4
    sliding_window_size: int = 32
    num_last_tokens: int = 64
    use_npu_optimization: bool = True      # kept for signature compatibility; no effect on B200
    npu_block_size: int = 64
    use_gradient_analysis: bool = False
    analysis_warmup_steps: int = 100

    def to_trianglemix_config(self) -> Optional[TriangleMixConfig]:
        if not self.enable_trianglemix:
            return None
        return TriangleMixConfig(
            num_sink_tokens=self.num_sink_tokens,
            sliding_window_size=self.sliding_window_size,
            num_last_tokens=self.num_last_tokens,
            num_triangle_layers=self.num_triangle_layers,
            triangle_layer_indices=self.triangle_layer_indices if self.triangle_layer_indices else None,
        )

    @staticmethod
    def for_ascend_npu(num_triangle_layers: int = 12,
                       enable_gradient_analysis: bool = True) -> "TriangleMixInferenceConfig":
        return TriangleMixInferenceConfig(
            enable_trianglemix=True,
            num_triangle_layers=num_triangle_layers,
            use_npu_optimization=True,
            use_gradient_analysis=enable_gradient_analysis,
        )

    @staticmethod
    def for_qwen3(model_size: str = "14B",
                  num_triangle_layers: Optional[int] = None) -> "TriangleMixInferenceConfig":
        if num_triangle_layers is None:
            num_triangle_layers = {"14B": 10, "32B": 12}.get(model_size, 8)
        return TriangleMixInferenceConfig(
            enable_trianglemix=True,
            num_triangle_layers=num_triangle_layers,
            num_sink_tokens=4,
            sliding_window_size=32,
            num_last_tokens=64,
            use_npu_optimization=True,
        )

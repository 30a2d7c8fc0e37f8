"""Producer for the analyzer's gradients (second half).

The reference ships `TriangleMixAnalyzer` (vllm/attention/trianglemix_attention.py:344-409): it records one
"Middle Q-K gradient" per layer and picks the layers with the LOWEST mean as triangle layers -- but nothing in
the reference produces those gradients.  This module does, on the GPU, from forward passes of the attention
kernels only (no autograd graph over N x N scores):

Put a gate g on the Middle Q-K block (the part of the causal mask a triangle layer drops,
`_create_middle_mask`, :133-171): softmax weights of middle keys are multiplied by g, g = 1 is dense attention,
g = 0 is triangle attention.  For query row i with triangle-part output O_tri, middle-part output O_mid and
middle softmax mass m = sigmoid(lse_mid - lse_tri):

    O_i(g) = ((1 - m) O_tri + g m O_mid) / ((1 - m) + g m)
    dO_i/dg at g = 1  =  m (1 - m) (O_mid - O_tri)
    dL/dg             =  sum_i  dO_i . m_i (1 - m_i) (O_mid,i - O_tri,i)        (dO = dL/dO, from the model's backward)

Two kernel launches per layer (mask pieces TRIANGLE and MIDDLE, each with its log-sum-exp) and an elementwise
reduction give the per-(token, head) gradient exactly.  `TriangleLayerSelector` feeds |gradient| into the
unchanged analyzer."""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import ops
from .trianglemix_attention import TriangleMixAnalyzer, TriangleMixConfig


def middle_qk_gradient(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, d_out: torch.Tensor, seq_lens: Sequence[int],
                       scale: float, si: int, sl: int, last: int, kernel: str = "auto") -> torch.Tensor:
    """dL/dg per (token, head) for a gate g on the Middle Q-K block, at g = 1 (see the module docstring).

    q, d_out [T, Hq, D], k / v [T, Hkv, D] bf16 on a CUDA device (varlen layout, sequences of `seq_lens` back
    to back); d_out = dL/d(attention output).  Returns float32 [T, Hq].  Rows without middle keys give 0."""
    t, hq, d = q.shape
    grad = torch.zeros((t, hq), dtype=torch.float32, device=q.device)
    start = 0
    for n in seq_lens:
        n = int(n)
        sl_ = slice(start, start + n)
        start += n
        if n == 0:
            continue
        cu = torch.tensor([0, n], dtype=torch.int32, device=q.device)
        o_tri, lse_tri = ops.prefill_attn(q[sl_], k[sl_], v[sl_], cu, n, scale, si, sl, last, ops.TRIANGLE,
                                          return_lse=True, kernel=kernel)
        # the middle piece alone: its packed bits select the keys, mode NONE touches every tile
        bits = ops.mask_bits(n, si, sl, last, "middle", q.device)
        o_mid, lse_mid = ops.prefill_attn(q[sl_], k[sl_], v[sl_], cu, n, scale, 0, 0, 0, ops.NONE, bitmask=bits,
                                          return_lse=True, kernel=kernel)
        m = torch.sigmoid(lse_mid - lse_tri).transpose(0, 1)            # [n, Hq]; lse_mid = -inf -> 0
        m = torch.nan_to_num(m, nan=0.0)
        diff = (o_mid.float() - o_tri.float()) * d_out[sl_].float()
        grad[sl_] = m * (1.0 - m) * diff.sum(dim=-1)
    return grad


class TriangleLayerSelector:
    """Accumulates Middle Q-K gradients per layer over calibration batches and hands the reference's analyzer
    (`TriangleMixAnalyzer.get_triangle_layers`, :375-405) what it expects: the layers whose Middle Q-K block
    matters least (lowest mean |dL/dg|) become triangle layers."""

    def __init__(self, num_layers: int, config: Optional[TriangleMixConfig] = None):
        self.config = config or TriangleMixConfig()
        self.analyzer = TriangleMixAnalyzer(num_layers)

    def observe(self, layer_idx: int, q, k, v, d_out, seq_lens: Sequence[int], scale: float, kernel: str = "auto"):
        si, sl, last = self.config.mask_params()
        g = middle_qk_gradient(q, k, v, d_out, seq_lens, scale, si, sl, last, kernel=kernel)
        self.analyzer.record_middle_qk_gradient(layer_idx, g.abs())
        return g

    def select(self, num_triangle_layers: int) -> List[int]:
        return self.analyzer.get_triangle_layers(num_triangle_layers)

    def to_config(self, num_triangle_layers: int) -> TriangleMixConfig:
        si, sl, last = self.config.mask_params()
        return TriangleMixConfig(si, sl, last, triangle_layer_indices=self.select(num_triangle_layers))
